// detector.cu — the reference's object-detector entry points over the B200 path.
//
//   tk_object_detector_create / _destroy / _detect / _update_thresholds
//       src/vision/tk_object_detector.h:96,102,138-143,166
//       src/vision/tk_object_detector.c:82-219 (lifecycle, detect, thresholds)
//
// detect() = preprocess_frame (:223-259) -> run_inference (:261-301) ->
// postprocess_detections (:303-368), all on the device: the frame is uploaded
// once, the fp32 CHW tensor is produced by the fused pre-process kernel, the
// network bound with tk_b200_object_detector_bind_inference reads it and writes
// the [P][5+C] head IN HBM (the reference round-trips both through host memory,
// :266-269,286-294), decode + class-wise NMS + coordinate restore run on the
// same stream, and only the kept records come back.  ONNX Runtime itself is a
// third-party engine outside this path; without a bound network detect()
// fails with TK_ERROR_INFERENCE_FAILED — there is no CPU fallback.
#include "tk_common.cuh"
#include <stdlib.h>
#include <time.h>

#include <stdlib.h>
#include <string.h>

struct tk_object_detector_s {
    tk_object_detector_config_t config;
    tk_b200_inference_fn infer;
    void* infer_user;
    uint32_t num_proposals;
    tk_b200_activation_e activation;
    cudaStream_t stream;
    // grow-only device buffers: 0 frames, 1 tensor, 2 logits, 3 candidates + detections + counts
    void* d_buf[4];
    size_t d_cap[4];
    uint8_t* h_pinned;      // detections + counts
    size_t h_cap;
};

namespace tkb {

constexpr uint32_t kDetCap = 4096;      // pre-NMS candidates per frame
constexpr uint32_t kDetMax = 500;       // TK_OBJECT_DETECTOR_MAX_DETECTIONS, src/vision/tk_object_detector.c:39

struct DevScope {
    int prev = -1; bool ok = false;
    explicit DevScope(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        ok = cudaSetDevice(dev) == cudaSuccess;
        if (!ok) cudaGetLastError();
    }
    ~DevScope() { if (prev >= 0) cudaSetDevice(prev); }
};

static tk_error_code_t reserve(tk_object_detector_s* d, int slot, size_t bytes) {
    if (d->d_cap[slot] >= bytes) return TK_SUCCESS;
    if (d->d_buf[slot]) { cudaStreamSynchronize(d->stream); cudaFree(d->d_buf[slot]); d->d_buf[slot] = nullptr; d->d_cap[slot] = 0; }
    const size_t want = bytes + bytes / 4 + 256;
    if (cudaMalloc(&d->d_buf[slot], want) != cudaSuccess) {
        tk_error_set_detail("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(cudaGetLastError()));
        return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
    }
    d->d_cap[slot] = want;
    return TK_SUCCESS;
}

static tk_error_code_t reserve_pinned(tk_object_detector_s* d, size_t bytes) {
    if (d->h_cap >= bytes) return TK_SUCCESS;
    if (d->h_pinned) { cudaStreamSynchronize(d->stream); cudaFreeHost(d->h_pinned); d->h_pinned = nullptr; d->h_cap = 0; }
    if (cudaHostAlloc(reinterpret_cast<void**>(&d->h_pinned), bytes, cudaHostAllocDefault) != cudaSuccess) {
        cudaGetLastError();
        return TK_ERROR_OUT_OF_MEMORY;
    }
    d->h_cap = bytes;
    return TK_SUCCESS;
}

static size_t frame_bytes(const tk_video_frame_t& f, uint32_t* out_stride) {
    uint32_t bpp;
    switch (f.format) {
        case TK_PIXEL_FORMAT_RGB8: case TK_PIXEL_FORMAT_BGR8: bpp = 3; break;
        case TK_PIXEL_FORMAT_RGBA8: bpp = 4; break;
        case TK_PIXEL_FORMAT_I420: bpp = 1; break;
        default: return 0;
    }
    const uint32_t stride = f.stride ? f.stride : f.width * bpp;
    if (stride < f.width * bpp) return 0;
    *out_stride = stride;
    size_t b = (size_t)stride * f.height;
    if (f.format == TK_PIXEL_FORMAT_I420) b += 2 * (size_t)(stride / 2) * (f.height / 2);
    return b;
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_object_detector_create(tk_object_detector_t** out, const tk_object_detector_config_t* config) {
    if (!out || !config || !config->model_path) return TK_ERROR_INVALID_ARGUMENT;      // tk_object_detector.c:84
    *out = nullptr;
    if (config->input_width == 0 || config->input_height == 0) return TK_ERROR_INVALID_ARGUMENT;
    if (config->backend != TK_VISION_BACKEND_CUDA) {
        tk_error_set_detail("libtk_b200 has no CPU / Metal / ROCm path: backend must be TK_VISION_BACKEND_CUDA");
        return TK_ERROR_NOT_IMPLEMENTED;
    }
    DevScope scope(config->gpu_device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    if (!device_info()) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    tk_object_detector_s* d = static_cast<tk_object_detector_s*>(calloc(1, sizeof(tk_object_detector_s)));
    if (!d) return TK_ERROR_OUT_OF_MEMORY;
    d->config = *config;
    if (cudaStreamCreateWithFlags(&d->stream, cudaStreamNonBlocking) != cudaSuccess) {
        cudaGetLastError(); free(d);
        return TK_ERROR_GPU_STREAM_CREATION_FAILED;
    }
    *out = d;
    return TK_SUCCESS;
}

extern "C" void tk_object_detector_destroy(tk_object_detector_t** pd) {
    if (!pd || !*pd) return;
    tk_object_detector_s* d = *pd;
    {
        DevScope scope(d->config.gpu_device_id);
        if (d->stream) cudaStreamSynchronize(d->stream);
        for (int i = 0; i < 4; ++i) if (d->d_buf[i]) cudaFree(d->d_buf[i]);
        if (d->h_pinned) cudaFreeHost(d->h_pinned);
        if (d->stream) cudaStreamDestroy(d->stream);
        cudaGetLastError();
    }
    free(d);
    *pd = nullptr;
}

// src/vision/tk_object_detector.c:212-219 (the caller serialises, tk_vision_pipeline.c:131-150)
extern "C" void tk_object_detector_update_thresholds(tk_object_detector_t* d, float confidence_threshold, float iou_threshold) {
    if (!d) return;
    d->config.confidence_threshold = confidence_threshold;
    d->config.iou_threshold = iou_threshold;
}

extern "C" tk_error_code_t tk_b200_object_detector_bind_inference(tk_object_detector_t* d, tk_b200_inference_fn fn, void* user,
                                                                  uint32_t num_proposals, tk_b200_activation_e activation) {
    if (!d || !fn || num_proposals == 0) return TK_ERROR_INVALID_ARGUMENT;
    if ((unsigned)activation > TK_B200_ACT_SOFTMAX) return TK_ERROR_INVALID_ARGUMENT;
    d->infer = fn; d->infer_user = user; d->num_proposals = num_proposals; d->activation = activation;
    return TK_SUCCESS;
}

extern "C" tk_error_code_t tk_b200_object_detector_detect_batch(tk_object_detector_t* d, const tk_video_frame_t* frames, uint32_t n,
                                                                tk_detection_result_t** out_results, size_t* out_counts) {
    if (!d || !frames || !out_results || !out_counts) return TK_ERROR_INVALID_ARGUMENT;      // tk_object_detector.c:189-191
    for (uint32_t i = 0; i < n; ++i) { out_results[i] = nullptr; out_counts[i] = 0; }
    if (n == 0) return TK_SUCCESS;
    if (!d->infer) {
        tk_error_set_detail("no detector network bound (tk_b200_object_detector_bind_inference)");
        return TK_ERROR_INFERENCE_FAILED;
    }
    uint32_t stride = 0;
    const tk_video_frame_t& f0 = frames[0];
    if (!f0.data || f0.width < 2 || f0.height < 2) return TK_ERROR_INVALID_ARGUMENT;
    const size_t fbytes = frame_bytes(f0, &stride);
    if (fbytes == 0) return TK_ERROR_INVALID_ARGUMENT;
    for (uint32_t i = 1; i < n; ++i) {
        const tk_video_frame_t& f = frames[i];
        if (!f.data || f.width != f0.width || f.height != f0.height || f.format != f0.format ||
            (f.stride ? f.stride : stride) != stride) {
            tk_error_set_detail("detect_batch: frame %u differs in geometry or format from frame 0", i);
            return TK_ERROR_INVALID_ARGUMENT;
        }
    }
    DevScope scope(d->config.gpu_device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;

    const uint32_t W = d->config.input_width, H = d->config.input_height;
    const uint32_t C = (uint32_t)d->config.class_count, P = d->num_proposals, L = 5 + C;
    const size_t pitch = (fbytes + 63) & ~(size_t)63;                  // 16-byte rule of the TMA row ring, +slack for its overrun
    const size_t tensor_bytes = (size_t)n * 3 * W * H * sizeof(float);
    const size_t logit_bytes = (size_t)n * P * L * sizeof(float);
    const size_t cand_bytes = (size_t)n * kDetCap * sizeof(tk_b200_candidate_t);
    const size_t det_bytes = (size_t)n * kDetMax * sizeof(tk_b200_candidate_t);
    const size_t off_det = (cand_bytes + 255) & ~(size_t)255, off_cnt = (off_det + det_bytes + 255) & ~(size_t)255;
    tk_error_code_t rc;
    if ((rc = reserve(d, 0, pitch * n + 64)) != TK_SUCCESS) return rc;
    if ((rc = reserve(d, 1, tensor_bytes)) != TK_SUCCESS) return rc;
    if ((rc = reserve(d, 2, logit_bytes)) != TK_SUCCESS) return rc;
    if ((rc = reserve(d, 3, off_cnt + 8 * (size_t)n + 256)) != TK_SUCCESS) return rc;
    if ((rc = reserve_pinned(d, off_cnt - off_det + 8 * (size_t)n)) != TK_SUCCESS) return rc;

    uint8_t* d_frames = static_cast<uint8_t*>(d->d_buf[0]);
    float* d_tensor = static_cast<float*>(d->d_buf[1]);
    float* d_logits = static_cast<float*>(d->d_buf[2]);
    uint8_t* base = static_cast<uint8_t*>(d->d_buf[3]);
    tk_b200_candidate_t* d_c = reinterpret_cast<tk_b200_candidate_t*>(base);
    tk_b200_candidate_t* d_d = reinterpret_cast<tk_b200_candidate_t*>(base + off_det);
    uint32_t* d_cnt = reinterpret_cast<uint32_t*>(base + off_cnt);     // [n] candidates, then [n] detections
    cudaStream_t st = d->stream;

    // TK_B200_DETECT_TRACE=1 (debug): synchronise after every phase and print where a call's time goes
    static const bool ptrace = [] { const char* e = getenv("TK_B200_DETECT_TRACE"); return e && atoi(e) != 0; }();
    static int ptrace_calls = 0;
    double pt[8] = {}; int pi = 0;
    auto mark = [&](bool sync) {
        if (!ptrace) return;
        if (sync) cudaStreamSynchronize(st);
        timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
        if (pi < 8) pt[pi++] = ts.tv_sec * 1e6 + ts.tv_nsec * 1e-3;
    };
    mark(false);
    // preprocess_frame, :223-259 (upload + kernel; constants of :248-249)
    for (uint32_t i = 0; i < n; ++i)
        TK_CUDA_TRY(cudaMemcpyAsync(d_frames + pitch * i, frames[i].data, fbytes, cudaMemcpyHostToDevice, st), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    mark(true);
    tk_b200_preprocess_batch_params_t pp{};
    pp.d_input = d_frames; pp.input_frame_pitch_bytes = pitch; pp.num_frames = n;
    pp.input_width = f0.width; pp.input_height = f0.height; pp.input_stride_bytes = stride; pp.input_format = f0.format;
    pp.d_output = d_tensor; pp.output_width = W; pp.output_height = H;
    pp.output_type = TK_B200_OUT_F32; pp.resize_mode = TK_B200_RESIZE_STRETCH; pp.arith = TK_B200_ARITH_EXACT;
    pp.mean = TkFloat3{0.485f, 0.456f, 0.406f};
    pp.std_dev = TkFloat3{0.229f, 0.224f, 0.225f};
    if ((rc = tk_b200_preprocess_batch(&pp, st)) != TK_SUCCESS) return rc;
    mark(true);

    // run_inference, :261-301 — the bound network, device pointers in and out
    const tk_error_code_t irc = d->infer(d->infer_user, d_tensor, n, W, H, d_logits, P, L, st);
    if (irc != TK_SUCCESS) {
        tk_error_set_detail("bound detector network returned %d", (int)irc);
        cudaStreamSynchronize(st);
        return TK_ERROR_INFERENCE_FAILED;
    }

    mark(true);
    // postprocess_detections, :303-368
    tk_b200_decode_params_t dp{};
    dp.d_logits = d_logits; dp.num_frames = n; dp.num_proposals = P; dp.num_classes = C;
    dp.confidence_threshold = d->config.confidence_threshold; dp.activation = d->activation;
    dp.d_candidates = d_c; dp.d_counts = d_cnt; dp.max_candidates = kDetCap;
    if ((rc = tk_kernels_decode_detections(&dp, st)) != TK_SUCCESS) return rc;
    mark(true);
    tk_b200_nms_params_t np{};
    np.d_candidates = d_c; np.d_counts = d_cnt; np.num_frames = n; np.max_candidates = kDetCap;
    np.iou_threshold = d->config.iou_threshold; np.max_detections = kDetMax; np.class_agnostic = 0;
    np.d_detections = d_d; np.d_num_detections = d_cnt + n;
    np.restore_src_width = f0.width; np.restore_src_height = f0.height; np.restore_in_width = W; np.restore_in_height = H;   // :365
    if ((rc = tk_kernels_nms_classwise(&np, st)) != TK_SUCCESS) return rc;
    mark(true);

    // detections and both count arrays are contiguous on the device (off_det .. off_cnt + 8 n): ONE copy brings them back
    const size_t back_bytes = off_cnt - off_det + 8 * (size_t)n;
    tk_b200_candidate_t* h_det = reinterpret_cast<tk_b200_candidate_t*>(d->h_pinned);
    const uint32_t* h_ncand = reinterpret_cast<const uint32_t*>(d->h_pinned + (off_cnt - off_det));   // [n] candidates
    uint32_t* h_ndet = reinterpret_cast<uint32_t*>(d->h_pinned + (off_cnt - off_det)) + n;           // [n] detections
    TK_CUDA_TRY(cudaMemcpyAsync(d->h_pinned, base + off_det, back_bytes, cudaMemcpyDeviceToHost, st), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    if (cudaStreamSynchronize(st) != cudaSuccess) {
        tk_error_set_detail("detect: %s", cudaGetErrorString(cudaGetLastError()));
        return TK_ERROR_INFERENCE_FAILED;
    }
    mark(false);
    if (ptrace && ++ptrace_calls % 50 == 0)
        fprintf(stderr, "detect trace (us, phases serialised): upload %.1f  pre-process %.1f  network %.1f  decode %.1f  nms %.1f  download+sync %.1f\n",
                pt[1] - pt[0], pt[2] - pt[1], pt[3] - pt[2], pt[4] - pt[3], pt[5] - pt[4], pt[6] - pt[5]);
    // The reference keeps every row that passes the threshold (:339-355); the kernels above cap a frame at kDetCap
    // candidates.  A frame that went over (a low threshold, a crowd) is done again without a cap — never truncated.
    bool redone = false;
    for (uint32_t i = 0; i < n; ++i) {
        if (h_ncand[i] <= kDetCap) continue;
        const uint32_t restore[4] = {f0.width, f0.height, W, H};
        if ((rc = postprocess_frame_uncapped(d_logits + (size_t)i * P * L, P, C, d->activation, d->config.confidence_threshold,
                                             d->config.iou_threshold, kDetMax, 0, restore, d_d + (size_t)i * kDetMax,
                                             d_cnt + n + i, st)) != TK_SUCCESS) {
            cudaStreamSynchronize(st);
            return rc;
        }
        TK_CUDA_TRY(cudaMemcpyAsync(h_ndet + i, d_cnt + n + i, 4, cudaMemcpyDeviceToHost, st), TK_ERROR_GPU_MEMORY_COPY_FAILED);
        TK_CUDA_TRY(cudaMemcpyAsync(h_det + (size_t)i * kDetMax, d_d + (size_t)i * kDetMax, kDetMax * sizeof(tk_b200_candidate_t),
                                    cudaMemcpyDeviceToHost, st), TK_ERROR_GPU_MEMORY_COPY_FAILED);
        redone = true;
    }
    if (redone && cudaStreamSynchronize(st) != cudaSuccess) {
        tk_error_set_detail("detect (uncapped frame): %s", cudaGetErrorString(cudaGetLastError()));
        return TK_ERROR_INFERENCE_FAILED;
    }

    // results array, :359-361 (label points into the caller's table, tk_object_detector.h:50-52)
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t k = h_ndet[i] < kDetMax ? h_ndet[i] : kDetMax;
        if (k == 0) continue;
        tk_detection_result_t* res = static_cast<tk_detection_result_t*>(calloc(k, sizeof(tk_detection_result_t)));
        if (!res) {
            for (uint32_t j = 0; j < i; ++j) { free(out_results[j]); out_results[j] = nullptr; out_counts[j] = 0; }
            return TK_ERROR_OUT_OF_MEMORY;
        }
        const tk_b200_candidate_t* src = h_det + (size_t)i * kDetMax;
        for (uint32_t j = 0; j < k; ++j) {
            res[j].class_id = src[j].class_id;
            res[j].label = (d->config.class_labels && src[j].class_id < d->config.class_count) ? d->config.class_labels[src[j].class_id] : nullptr;
            res[j].confidence = src[j].confidence;
            res[j].bbox = src[j].bbox;
        }
        out_results[i] = res; out_counts[i] = k;
    }
    return TK_SUCCESS;
}

extern "C" tk_error_code_t tk_object_detector_detect(tk_object_detector_t* detector, const tk_video_frame_t* video_frame,
                                                     tk_detection_result_t** out_results, size_t* out_result_count) {
    if (!detector || !video_frame || !out_results || !out_result_count) return TK_ERROR_INVALID_ARGUMENT;   // :189-191
    return tk_b200_object_detector_detect_batch(detector, video_frame, 1, out_results, out_result_count);
}
