/*
 * tk_shard_test.c — the multi-device harness from plain C: ONE process drives every visible B200 (or, on a one-GPU
 * box, several pipelines on device 0) through tk_b200_shard_process_host and compares every output byte with what a
 * single tk_b200_pipeline on device 0 gives for the same batch.  The reference has one device per dispatcher
 * (src/gpu/cuda/tk_cuda_dispatch.h:47-49); this is the form SURVEY §8e asks for — one host thread + one pipeline per
 * device, contiguous frame blocks, no collective, host gather by prefix sum.
 *
 * usage: tk_shard_test <frames.u8> <logits.f32> <num_frames> <width> <height> <num_proposals> <num_classes> [min_devices]
 * Prints "devices D frames N detections T checksum X e2e_ms_single A e2e_ms_sharded B"; exit 0 = identical.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "tk_b200_vision.h"

static double now_ms(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static int slurp_into(const char* path, void* dst, size_t want) {
    FILE* f = fopen(path, "rb");
    if (!f) return 0;
    size_t got = fread(dst, 1, want, f);
    fclose(f);
    return got == want;
}

#define CHECK(call) do { tk_error_code_t rc__ = (call); if (rc__ != TK_SUCCESS) { \
    fprintf(stderr, "%s -> %d (%s) %s\n", #call, (int)rc__, tk_error_to_string(rc__), tk_error_get_detail()); return 4; } } while (0)

int main(int argc, char** argv) {
    if (argc < 8) { fprintf(stderr, "usage: %s frames.u8 logits.f32 N W H P C [min_devices]\n", argv[0]); return 2; }
    const uint32_t N = (uint32_t)atoi(argv[3]), W = (uint32_t)atoi(argv[4]), H = (uint32_t)atoi(argv[5]);
    const uint32_t P = (uint32_t)atoi(argv[6]), C = (uint32_t)atoi(argv[7]);
    const int min_dev = argc > 8 ? atoi(argv[8]) : 1;
    const uint32_t TW = 96, TH = 64, MAXDET = 300;
    const size_t fbytes = (size_t)W * H * 3, lbytes = (size_t)P * (5 + C) * sizeof(float), tbytes = (size_t)TW * TH * 3 * sizeof(float);

    int ndev = tk_b200_device_count();
    if (ndev < 1) { fprintf(stderr, "no CUDA device\n"); return 3; }
    if (ndev < min_dev) { fprintf(stderr, "%d devices visible, %d wanted\n", ndev, min_dev); return 3; }
    int ids[64];
    uint32_t G = 0;
    if (ndev > 1) { for (int i = 0; i < ndev && i < 64; ++i) ids[G++] = i; }
    else { ids[0] = ids[1] = ids[2] = 0; G = 3; }                   /* one GPU: three pipelines on it */

    void *h_frames, *h_logits, *h_t1, *h_t2, *h_d1, *h_d2, *h_c1, *h_c2;
    CHECK(tk_b200_host_alloc(&h_frames, N * fbytes));
    CHECK(tk_b200_host_alloc(&h_logits, N * lbytes));
    CHECK(tk_b200_host_alloc(&h_t1, N * tbytes));
    CHECK(tk_b200_host_alloc(&h_t2, N * tbytes));
    CHECK(tk_b200_host_alloc(&h_d1, (size_t)N * MAXDET * sizeof(tk_b200_candidate_t)));
    CHECK(tk_b200_host_alloc(&h_d2, (size_t)N * MAXDET * sizeof(tk_b200_candidate_t)));
    CHECK(tk_b200_host_alloc(&h_c1, N * sizeof(uint32_t)));
    CHECK(tk_b200_host_alloc(&h_c2, N * sizeof(uint32_t)));
    if (!slurp_into(argv[1], h_frames, N * fbytes) || !slurp_into(argv[2], h_logits, N * lbytes)) { fprintf(stderr, "cannot read inputs\n"); return 2; }

    tk_b200_pipeline_config_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.device_id = 0; cfg.max_batch = N; cfg.chunk_frames = 4;
    cfg.input_width = W; cfg.input_height = H; cfg.input_format = TK_PIXEL_FORMAT_RGB8;
    cfg.tensor_width = TW; cfg.tensor_height = TH; cfg.tensor_type = TK_B200_OUT_F32;
    cfg.resize_mode = TK_B200_RESIZE_STRETCH; cfg.arith = TK_B200_ARITH_EXACT;
    cfg.mean = (TkFloat3){0.485f, 0.456f, 0.406f}; cfg.std_dev = (TkFloat3){0.229f, 0.224f, 0.225f};   /* tk_object_detector.c:248-249 */
    cfg.num_proposals = P; cfg.num_classes = C; cfg.activation = TK_B200_ACT_RAW;
    cfg.confidence_threshold = 0.5f; cfg.iou_threshold = 0.5f; cfg.max_detections = MAXDET;

    /* single device, the whole batch */
    tk_b200_pipeline_t* pipe = NULL;
    CHECK(tk_b200_pipeline_create(&pipe, &cfg));
    CHECK(tk_b200_pipeline_process_host(pipe, h_frames, (const float*)h_logits, N, h_t1, (tk_b200_candidate_t*)h_d1, (uint32_t*)h_c1));
    double t0 = now_ms();
    for (int r = 0; r < 3; ++r)
        CHECK(tk_b200_pipeline_process_host(pipe, h_frames, (const float*)h_logits, N, h_t1, (tk_b200_candidate_t*)h_d1, (uint32_t*)h_c1));
    const double ms_single = (now_ms() - t0) / 3;
    tk_b200_pipeline_destroy(&pipe);

    /* every device, one block each, from this one process */
    tk_b200_shard_t* sh = NULL;
    CHECK(tk_b200_shard_create(&sh, ids, G, &cfg));
    memset(h_t2, 0xAB, N * tbytes); memset(h_d2, 0xCD, (size_t)N * MAXDET * sizeof(tk_b200_candidate_t)); memset(h_c2, 0xEF, N * sizeof(uint32_t));
    CHECK(tk_b200_shard_process_host(sh, h_frames, (const float*)h_logits, N, h_t2, (tk_b200_candidate_t*)h_d2, (uint32_t*)h_c2));
    t0 = now_ms();
    for (int r = 0; r < 3; ++r)
        CHECK(tk_b200_shard_process_host(sh, h_frames, (const float*)h_logits, N, h_t2, (tk_b200_candidate_t*)h_d2, (uint32_t*)h_c2));
    const double ms_shard = (now_ms() - t0) / 3;

    int bad = 0;
    if (memcmp(h_t1, h_t2, N * tbytes)) { fprintf(stderr, "tensors differ\n"); bad = 1; }
    if (memcmp(h_c1, h_c2, N * sizeof(uint32_t))) { fprintf(stderr, "counts differ\n"); bad = 1; }
    const uint32_t* c1 = (const uint32_t*)h_c1;
    unsigned long long total = 0, sum = 1469598103934665603ull;
    for (uint32_t f = 0; f < N && !bad; ++f) {
        const uint32_t k = c1[f] < MAXDET ? c1[f] : MAXDET;
        const uint8_t* a = (const uint8_t*)h_d1 + (size_t)f * MAXDET * sizeof(tk_b200_candidate_t);
        const uint8_t* b = (const uint8_t*)h_d2 + (size_t)f * MAXDET * sizeof(tk_b200_candidate_t);
        if (memcmp(a, b, (size_t)k * sizeof(tk_b200_candidate_t))) { fprintf(stderr, "detections of frame %u differ\n", f); bad = 1; }
        for (size_t i = 0; i < (size_t)k * sizeof(tk_b200_candidate_t); ++i) sum = (sum ^ a[i]) * 1099511628211ull;
        total += k;
    }
    /* host gather: flat list + offsets */
    uint64_t* offs = (uint64_t*)malloc((N + 1) * sizeof(uint64_t));
    CHECK(tk_b200_shard_gather(NULL, (const uint32_t*)h_c2, N, MAXDET, NULL, offs));
    if (offs[N] != total) { fprintf(stderr, "gather total %llu != %llu\n", (unsigned long long)offs[N], total); bad = 1; }
    tk_b200_candidate_t* flat = (tk_b200_candidate_t*)malloc((size_t)(total ? total : 1) * sizeof(tk_b200_candidate_t));
    CHECK(tk_b200_shard_gather((const tk_b200_candidate_t*)h_d2, (const uint32_t*)h_c2, N, MAXDET, flat, offs));
    for (uint32_t f = 0; f < N && !bad; ++f)
        if (memcmp(flat + offs[f], (const tk_b200_candidate_t*)h_d2 + (size_t)f * MAXDET, (size_t)(offs[f + 1] - offs[f]) * sizeof(tk_b200_candidate_t))) {
            fprintf(stderr, "gathered frame %u differs\n", f); bad = 1;
        }
    printf("devices %u (visible %d) frames %u detections %llu checksum %016llx e2e_ms_single %.3f e2e_ms_sharded %.3f\n",
           G, ndev, N, total, sum, ms_single, ms_shard);
    free(flat); free(offs);
    tk_b200_shard_destroy(&sh);
    tk_b200_host_free(h_frames); tk_b200_host_free(h_logits); tk_b200_host_free(h_t1); tk_b200_host_free(h_t2);
    tk_b200_host_free(h_d1); tk_b200_host_free(h_d2); tk_b200_host_free(h_c1); tk_b200_host_free(h_c2);
    return bad;
}
