// decode_row.cuh — per-row decode shared by decode_kernel (decode.cu) and the fused post-process kernel (post.cu).
// Completes the candidate loop of postprocess_detections (src/vision/tk_object_detector.c:343-355).
#pragma once
#include "tk_common.cuh"

namespace tkb {

#ifndef TK_DEC_CLUSTER
#define TK_DEC_CLUSTER 8
#endif
static constexpr int kDecCluster = TK_DEC_CLUSTER;    // CTAs per frame in decode_kernel (experiment builds override it)
static constexpr int kDecThreads = 512;
static constexpr int kDecWarps = kDecThreads / 32;

struct DecK {
    const float* logits;
    const float* obj;          // PACKED: dense objectness [frames][P] (the logits may then sit in host memory; only
                               // the surviving rows are read from them)
    const uint32_t* bitmap;    // pass 1 done elsewhere (the pre-process kernel's scanner warps): [frames][nwf] ballot words
    uint32_t nwf;              // bitmap words per frame = ceil(P / 32)
    tk_b200_candidate_t* cands;
    uint32_t* counts;
    uint32_t P, C, L;          // proposals, classes, row length 5+C
    uint32_t rows_per_cta;     // multiple of 32
    uint32_t nwords;           // rows_per_cta / 32
    uint32_t cap;
    float thr;
    TraceBuf trace;            // debug: per-CTA phase stamps (tk_b200_debug_trace), buf == nullptr when off
};

template <int ACT>
__device__ __forceinline__ float act_fn(float x) {
    if constexpr (ACT == TK_B200_ACT_RAW) return x;
    else return tk_dm_sigmoidf(x);
}

static constexpr uint32_t kNone = 0xffffffffu;

// Objectness load of pass 1: 4 bytes out of every 340-byte row.  The L2::64B hint
// halves the DRAM fetch per row (the default fetches 128 B): 33 -> 25 us for
// 64 x 25200 rows (tools/exp_stride.cu).
__device__ __forceinline__ float ld_obj(const float* p) {
    float v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}

static constexpr uint32_t kDecStage = 256;     // candidate records a CTA stages in shared memory (7 KB)
static constexpr int kDecUnroll = 8;           // objectness loads in flight per thread

// One row's worth of loads: up to 128 class logits (lane + 32u) and the 5-word head (lanes 0..4: cx, cy, w, h,
// objectness), issued together — one memory round trip per row instead of one per 32 classes.  (Deeper prefetch —
// the next row's loads in flight, or four rows staged with cp.async — measured no faster: with 64 warps per SM in pass 2
// at once the pass is bound by instruction issue, not by the loads.)
struct RowRegs { float v4[4]; float head; };
// GLD: the row is in global memory (read-only path); otherwise plain loads (a row staged in shared memory).
template <bool GLD>
__device__ __forceinline__ float row_ld(const float* p) { if constexpr (GLD) return __ldg(p); else return *p; }
template <bool GLD = true>
__device__ __forceinline__ void load_row(const DecK& k, const float* row, uint32_t lane, RowRegs& q) {
#pragma unroll
    for (int u = 0; u < 4; ++u) q.v4[u] = (lane + 32u * u < k.C) ? row_ld<GLD>(row + 5 + lane + 32u * u) : 0.0f;
    q.head = lane < 5 ? row_ld<GLD>(row + lane) : 0.0f;
}

template <int ACT, bool GLD = true>
__device__ __forceinline__ tk_b200_candidate_t decode_row(const DecK& k, const float* row, uint32_t r, uint32_t lane, float* my_e,
                                                          const RowRegs& q) {
    // First-max class over raw logits: the reference's sequential scan
    //   best = 0; for j >= 1: if (v[j] > v[best]) best = j;      (tie rule of tk_attribute_classifier.c:83-88)
    // as two warp reductions on order-preserving integer keys.  -0 and +0 compare equal
    // (canonicalised by + 0.0f), a NaN at j >= 1 never wins (key 0, below every number),
    // a NaN at j == 0 is never displaced (key 0xffffffff); ties -> lowest index.
    uint32_t lk = 0, li = kNone; float lv = 0.0f;
    auto consider = [&](uint32_t j, float v) {
        const uint32_t key = (v != v) ? (j == 0 ? 0xffffffffu : 0u) : tk_dm_float_order_key(__fadd_rn(v, 0.0f));
        if (li == kNone || key > lk) { lk = key; li = j; lv = v; }
    };
#pragma unroll
    for (int u = 0; u < 4; ++u)
        if (lane + 32u * u < k.C) consider(lane + 32u * u, q.v4[u]);
    for (uint32_t j = lane + 128; j < k.C; j += 32) consider(j, row_ld<GLD>(row + 5 + j));
    const uint32_t mk = __reduce_max_sync(0xffffffffu, lk);
    const uint32_t bidx = __reduce_min_sync(0xffffffffu, (li != kNone && lk == mk) ? li : kNone);
    const uint32_t best = (bidx == kNone) ? 0u : bidx;
    const float bestv = __shfl_sync(0xffffffffu, lv, best & 31u);  // lane (best mod 32) owns class `best`
    float cls = 0.0f;
    if (k.C == 0) cls = 1.0f;
    else if constexpr (ACT == TK_B200_ACT_RAW) cls = bestv;
    else if constexpr (ACT == TK_B200_ACT_SIGMOID) cls = tk_dm_sigmoidf(bestv);
    else {
        // class softmax in the arithmetic order of tests/tk_gpu_softmax_test.cpp:31-49:
        // exponentials in parallel, the sum strictly in index order.
        for (uint32_t j = lane; j < k.C; j += 32) my_e[j] = tk_dm_expf(__fsub_rn(row_ld<GLD>(row + 5 + j), bestv));
        __syncwarp();
        float sum = 0.0f;
        if (lane == 0) {
            for (uint32_t j = 0; j < k.C; ++j) sum = __fadd_rn(sum, my_e[j]);
            cls = __fdiv_rn(tk_dm_expf(__fsub_rn(bestv, bestv)), sum);
        }
        __syncwarp();
    }
    tk_b200_candidate_t c;
    const float cx = __shfl_sync(0xffffffffu, q.head, 0), cy = __shfl_sync(0xffffffffu, q.head, 1);
    const float bw = __shfl_sync(0xffffffffu, q.head, 2), bh = __shfl_sync(0xffffffffu, q.head, 3);
    const float obj = act_fn<ACT>(__shfl_sync(0xffffffffu, q.head, 4));
    c.index = (int32_t)r;
    c.bbox.x = __float2int_rz(__fsub_rn(cx, __fmul_rn(bw, 0.5f)));   // :347  (int)(cx - w/2)
    c.bbox.y = __float2int_rz(__fsub_rn(cy, __fmul_rn(bh, 0.5f)));   // :348
    c.bbox.w = __float2int_rz(bw);                                    // :349
    c.bbox.h = __float2int_rz(bh);                                    // :350
    c.confidence = tk_dm_canon_nan(__fmul_rn(obj, cls));             // valid on lane 0 (cls)
    c.class_id = best;
    return c;
}


}  // namespace tkb
