/*
 * tk_detector_flow_test.c — the reference's detector call sequence, in the reference's own language (C),
 * against libtk_b200.so.  Mirrors what tk_vision_pipeline_process_frame does around the detector
 * (src/vision/tk_vision_pipeline.c:439: tk_object_detector_detect, then free_results) and the frame
 * of tests/tk_vision_full_test.c:30-32 (a 640x480 RGB8 frame; zero-filled there, any bytes here).
 *
 * usage: tk_detector_flow_test <logits.f32> <num_proposals> <num_classes> <frame.rgb> <width> <height>
 * The detector network is a stand-in: a device-to-device copy of the head tensor read from
 * <logits.f32> (ONNX Runtime and the model are not part of the reference tree).  Prints one line per
 * detection: class_id label confidence-bits x y w h.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cuda_runtime.h>
#include "tk_b200_vision.h"

typedef struct { float* d_logits; size_t bytes; int calls; } fake_net_t;

static tk_error_code_t fake_network(void* user, const float* d_input, uint32_t num_frames, uint32_t in_w, uint32_t in_h,
                                    float* d_output, uint32_t num_proposals, uint32_t row_length, cudaStream_t stream) {
    fake_net_t* net = (fake_net_t*)user;
    (void)d_input; (void)in_w; (void)in_h;
    size_t need = (size_t)num_frames * num_proposals * row_length * sizeof(float);
    if (need > net->bytes) return TK_ERROR_INVALID_OUTPUT_TENSOR;
    net->calls++;
    return cudaMemcpyAsync(d_output, net->d_logits, need, cudaMemcpyDeviceToDevice, stream) == cudaSuccess
               ? TK_SUCCESS : TK_ERROR_INFERENCE_FAILED;
}

static void* slurp(const char* path, size_t want) {
    FILE* f = fopen(path, "rb");
    if (!f) return NULL;
    void* p = malloc(want);
    size_t got = p ? fread(p, 1, want, f) : 0;
    fclose(f);
    if (got != want) { free(p); return NULL; }
    return p;
}

int main(int argc, char** argv) {
    if (argc != 7) { fprintf(stderr, "usage: %s logits.f32 P C frame.rgb W H\n", argv[0]); return 2; }
    const uint32_t P = (uint32_t)atoi(argv[2]), C = (uint32_t)atoi(argv[3]);
    const uint32_t W = (uint32_t)atoi(argv[5]), H = (uint32_t)atoi(argv[6]);
    const size_t lbytes = (size_t)P * (5 + C) * sizeof(float);
    float* h_logits = (float*)slurp(argv[1], lbytes);
    uint8_t* h_frame = (uint8_t*)slurp(argv[4], (size_t)W * H * 3);
    if (!h_logits || !h_frame) { fprintf(stderr, "cannot read inputs\n"); return 2; }

    static char names[128][16];
    static const char* labels[128];
    for (uint32_t i = 0; i < C && i < 128; ++i) { snprintf(names[i], sizeof names[i], "class%u", i); labels[i] = names[i]; }

    fake_net_t net = {0};
    net.bytes = lbytes;
    if (cudaMalloc((void**)&net.d_logits, lbytes) != cudaSuccess ||
        cudaMemcpy(net.d_logits, h_logits, lbytes, cudaMemcpyHostToDevice) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 3; }

    tk_object_detector_config_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.backend = TK_VISION_BACKEND_CUDA;
    cfg.gpu_device_id = 0;
    cfg.model_path = (tk_path_t*)"yolov5nu.onnx";            /* opaque to the library */
    cfg.input_width = 640; cfg.input_height = 640;          /* src/vision/tk_vision_pipeline.c:368-378 */
    cfg.class_labels = labels; cfg.class_count = C;
    cfg.confidence_threshold = 0.5f; cfg.iou_threshold = 0.5f;

    tk_object_detector_t* det = NULL;
    tk_error_code_t rc = tk_object_detector_create(&det, &cfg);
    if (rc != TK_SUCCESS) { fprintf(stderr, "create: %d %s\n", (int)rc, tk_error_get_detail()); return 4; }
    rc = tk_b200_object_detector_bind_inference(det, fake_network, &net, P, TK_B200_ACT_RAW);
    if (rc != TK_SUCCESS) { fprintf(stderr, "bind: %d\n", (int)rc); return 4; }

    tk_video_frame_t frame;
    memset(&frame, 0, sizeof frame);
    frame.width = W; frame.height = H; frame.stride = W * 3; frame.format = TK_PIXEL_FORMAT_RGB8; frame.data = h_frame;

    for (int pass = 0; pass < 2; ++pass) {                  /* twice: buffers are reused across calls */
        tk_detection_result_t* results = NULL;
        size_t count = 0;
        rc = tk_object_detector_detect(det, &frame, &results, &count);
        if (rc != TK_SUCCESS) { fprintf(stderr, "detect: %d %s\n", (int)rc, tk_error_get_detail()); return 5; }
        if (pass == 1)
            for (size_t i = 0; i < count; ++i) {
                uint32_t bits; memcpy(&bits, &results[i].confidence, 4);
                printf("%u %s %08x %d %d %d %d\n", results[i].class_id, results[i].label ? results[i].label : "-", bits,
                       results[i].bbox.x, results[i].bbox.y, results[i].bbox.w, results[i].bbox.h);
            }
        tk_object_detector_free_results(&results);
        if (results != NULL) return 6;
    }
    tk_object_detector_update_thresholds(det, 0.9f, 0.5f);
    tk_object_detector_destroy(&det);
    if (det != NULL || net.calls != 2) return 7;
    cudaFree(net.d_logits); free(h_logits); free(h_frame);
    return 0;
}
