// decode.cu — YOLO-head decode + confidence threshold + ORDERED stream compaction.
//
// Completes  postprocess_detections' candidate loop
//              (src/vision/tk_object_detector.c:343-355: row pointer :344, strict
//               `row[4] > thr` :345, C-truncated box :347-350, class/confidence
//               selection elided at :351) and replaces the atomicAdd compaction of
//            tk_cuda_nms_kernel (src/gpu/extensions/cuda/tk_cuda_tensor_ops.cu:452-482),
//            whose output order is non-deterministic.
//
// B200 design (HBM-bound scan, no tensor cores):
//   * One thread-block CLUSTER of 8 CTAs per frame; CTA r scans a contiguous
//     slice of the frame's proposal rows.
//   * Pass 1 reads only the objectness word of each row (one 32-byte sector out
//     of the 340-byte row), thresholds it and records a warp-ballot bitmap in
//     shared memory.  Rows that fail — almost all of them — are never read again.
//   * The 8 CTA totals are exchanged through distributed shared memory, giving
//     every CTA its exclusive offset inside the frame: the output is in
//     ascending proposal order with no atomics and no second launch.
//   * Pass 2: one warp per surviving row loads the whole row coalesced, finds
//     the first-max class by shuffle reduction, and writes the 28-byte record.
//   * Sigmoid / softmax use tk_dm_expf (include/tk_b200_detmath.h), the same
//     source the oracle compiles, so thresholds and scores agree bit for bit.
#include "tk_common.cuh"

#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace tkb {

static constexpr int kDecCluster = 8;
static constexpr int kDecThreads = 512;
static constexpr int kDecWarps = kDecThreads / 32;

struct DecK {
    const float* logits;
    tk_b200_candidate_t* cands;
    uint32_t* counts;
    uint32_t P, C, L;          // proposals, classes, row length 5+C
    uint32_t rows_per_cta;     // multiple of 32
    uint32_t nwords;           // rows_per_cta / 32
    uint32_t cap;
    float thr;
};

template <int ACT>
__device__ __forceinline__ float act_fn(float x) {
    if constexpr (ACT == TK_B200_ACT_RAW) return x;
    else return tk_dm_sigmoidf(x);
}

// (value, index) pairs for the reference's sequential first-max scan
//   best = 0; for j >= 1: if (v[j] > v[best]) best = j;
// i == kNone means "no element".  Only element 0 can carry a NaN value (NaNs at
// j >= 1 never win a `>` and are dropped when read); a NaN at element 0 is never
// displaced, exactly as in the sequential scan.
static constexpr uint32_t kNone = 0xffffffffu;
__device__ __forceinline__ void amax_combine(float& v, uint32_t& i, float ov, uint32_t oi) {
    if (oi == kNone) return;
    if (i == kNone) { v = ov; i = oi; return; }
    if (v != v) return;                       // NaN at element 0 stays
    if (ov != ov) { v = ov; i = oi; return; }
    if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
}

template <int ACT>
__global__ void __launch_bounds__(kDecThreads) decode_kernel(const DecK k) {
    extern __shared__ __align__(16) uint8_t dsm[];
    uint32_t* flags = reinterpret_cast<uint32_t*>(dsm);            // [nwords] ballot bitmap
    uint32_t* wpre = flags + k.nwords;                             // [nwords] exclusive prefix of popc
    float* ebuf = reinterpret_cast<float*>(wpre + k.nwords);       // softmax only: [warps][C]
    __shared__ uint32_t cta_tot[kDecCluster];                      // filled by every CTA of the cluster

    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t rank = cluster.block_rank();
    const uint32_t frame = blockIdx.x / kDecCluster;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* fl = k.logits + (size_t)frame * k.P * k.L;
    const uint32_t row0 = rank * k.rows_per_cta;

    // ---- pass 1: objectness threshold, ballot bitmap ------------------------
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");   // "I am resident" (waited on before DSMEM)
    for (uint32_t base = 0; base < k.rows_per_cta; base += 4 * kDecThreads) {
        float obj[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint32_t r = row0 + base + u * kDecThreads + tid;
            uint32_t lr = base + u * kDecThreads + tid;
            obj[u] = (lr < k.rows_per_cta && r < k.P) ? __ldg(fl + (size_t)r * k.L + 4) : __int_as_float(0x7fc00000);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint32_t lr = base + u * kDecThreads + tid;
            bool pass = act_fn<ACT>(obj[u]) > k.thr;               // NaN never passes
            uint32_t b = __ballot_sync(0xffffffffu, pass);
            if (lane == 0 && lr < k.rows_per_cta) flags[lr >> 5] = b;
        }
    }
    __syncthreads();
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");     // every peer CTA has started

    // ---- exclusive prefix over the bitmap words + cluster exchange ----------
    if (warp == 0) {
        uint32_t run = 0;
        for (uint32_t w0 = 0; w0 < k.nwords; w0 += 32) {
            uint32_t w = w0 + lane;
            uint32_t c = w < k.nwords ? __popc(flags[w]) : 0u;
            uint32_t inc = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
                if ((int)lane >= d) inc += t;
            }
            if (w < k.nwords) wpre[w] = run + inc - c;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane < kDecCluster)
            *cluster.map_shared_rank(&cta_tot[rank], lane) = run;  // DSMEM store into every peer
    }
    cluster.sync();
    uint32_t frame_pre = 0, frame_tot = 0;
#pragma unroll
    for (int r = 0; r < kDecCluster; ++r) {
        uint32_t t = cta_tot[r];
        frame_pre += (uint32_t)r < rank ? t : 0u;
        frame_tot += t;
    }
    if (rank == 0 && tid == 0) k.counts[frame] = frame_tot;

    // ---- pass 2: one warp per surviving row ---------------------------------
    tk_b200_candidate_t* out = k.cands + (size_t)frame * k.cap;
    float* my_e = ebuf + (size_t)warp * k.C;
    for (uint32_t w = warp; w < k.nwords; w += kDecWarps) {
        const uint32_t word = flags[w];
        uint32_t bits = word;
        while (bits) {
            const uint32_t b = __ffs(bits) - 1;
            bits &= bits - 1;
            const uint32_t pos = frame_pre + wpre[w] + __popc(word & ((1u << b) - 1u));
            if (pos >= k.cap) { bits = 0; break; }                 // later rows only have larger pos
            const uint32_t r = row0 + w * 32 + b;
            const float* row = fl + (size_t)r * k.L;
            // first-max class over raw logits (tie rule of tk_attribute_classifier.c:83-88)
            float bv = 0.0f; uint32_t bi = kNone;
            for (uint32_t j = lane; j < k.C; j += 32) {
                float v = __ldg(row + 5 + j);
                if (j == 0 || v == v) amax_combine(bv, bi, v, j);
            }
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                float ov = __shfl_xor_sync(0xffffffffu, bv, d);
                uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, d);
                amax_combine(bv, bi, ov, oi);
            }
            const uint32_t best = (bi == kNone) ? 0u : bi;
            const float bestv = bv;
            float cls = 0.0f;
            if (k.C == 0) cls = 1.0f;
            else if constexpr (ACT == TK_B200_ACT_RAW) cls = bestv;
            else if constexpr (ACT == TK_B200_ACT_SIGMOID) cls = tk_dm_sigmoidf(bestv);
            else {
                // class softmax in the arithmetic order of tests/tk_gpu_softmax_test.cpp:31-49:
                // exponentials in parallel, the sum strictly in index order.
                for (uint32_t j = lane; j < k.C; j += 32) my_e[j] = tk_dm_expf(__fsub_rn(__ldg(row + 5 + j), bestv));
                __syncwarp();
                float sum = 0.0f;
                if (lane == 0) {
                    for (uint32_t j = 0; j < k.C; ++j) sum = __fadd_rn(sum, my_e[j]);
                    cls = __fdiv_rn(tk_dm_expf(__fsub_rn(bestv, bestv)), sum);
                }
                __syncwarp();
            }
            if (lane == 0) {
                float cx = __ldg(row + 0), cy = __ldg(row + 1), bw = __ldg(row + 2), bh = __ldg(row + 3);
                float obj = act_fn<ACT>(__ldg(row + 4));
                tk_b200_candidate_t c;
                c.index = (int32_t)r;
                c.bbox.x = __float2int_rz(__fsub_rn(cx, __fmul_rn(bw, 0.5f)));   // :347  (int)(cx - w/2)
                c.bbox.y = __float2int_rz(__fsub_rn(cy, __fmul_rn(bh, 0.5f)));   // :348
                c.bbox.w = __float2int_rz(bw);                                    // :349
                c.bbox.h = __float2int_rz(bh);                                    // :350
                c.confidence = tk_dm_canon_nan(__fmul_rn(obj, cls));
                c.class_id = best;
                out[pos] = c;
            }
        }
    }
}

template <int ACT>
static tk_error_code_t launch_decode(const DecK& k, uint32_t frames, size_t smem, cudaStream_t stream) {
    auto kern = decode_kernel<ACT>;
    if (smem > 48 * 1024)
        TK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), TK_ERROR_GPU_CUDA_ERROR);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(frames * kDecCluster);
    cfg.blockDim = dim3(kDecThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = kDecCluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    TK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, k), TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE);
    return launch_status("decode_kernel");
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_kernels_decode_detections(const tk_b200_decode_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_logits || !p->d_counts || (!p->d_candidates && p->max_candidates)) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_frames == 0) return TK_SUCCESS;
    if (p->num_proposals == 0) return TK_ERROR_INVALID_DIMENSIONS;      // dims check of tk_object_detector.c:325-336
    if ((unsigned)p->activation > TK_B200_ACT_SOFTMAX) return TK_ERROR_INVALID_ARGUMENT;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    DecK k{};
    k.logits = p->d_logits; k.cands = p->d_candidates; k.counts = p->d_counts;
    k.P = p->num_proposals; k.C = p->num_classes; k.L = 5 + p->num_classes;
    uint32_t per = (k.P + kDecCluster - 1) / kDecCluster;
    k.rows_per_cta = (per + 31) & ~31u;
    k.nwords = k.rows_per_cta / 32;
    k.cap = p->max_candidates;
    k.thr = p->confidence_threshold;
    size_t smem = (size_t)k.nwords * 8 + (p->activation == TK_B200_ACT_SOFTMAX ? (size_t)kDecWarps * k.C * 4 : 0);
    if (smem > (size_t)dev->max_smem_optin - 1024) {
        tk_error_set_detail("decode: %u proposals x %u classes need %zu B of shared memory", k.P, k.C, smem);
        return TK_ERROR_NOT_IMPLEMENTED;
    }
    switch (p->activation) {
        case TK_B200_ACT_RAW:     return launch_decode<TK_B200_ACT_RAW>(k, p->num_frames, smem, stream);
        case TK_B200_ACT_SIGMOID: return launch_decode<TK_B200_ACT_SIGMOID>(k, p->num_frames, smem, stream);
        default:                  return launch_decode<TK_B200_ACT_SOFTMAX>(k, p->num_frames, smem, stream);
    }
}
