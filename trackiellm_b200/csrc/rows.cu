// rows.cu — row softmax with fused first-max argmax.
//
// Replaces  tk_cuda_softmax_kernel / tk_kernels_softmax
//             (src/gpu/cuda/tk_cuda_kernels.cu:312-393): one block per row with
//             blockDim == num_cols and a halving tree that is only correct for
//             power-of-two num_cols (:331-336) and rejects num_cols > 1024 (:369).
// Parity target: softmax_cpu of tests/tk_gpu_softmax_test.cpp:24-51 — max :33-38,
//             sum of exp(x-max) in INDEX ORDER :41-44, divide :47-49.
//
// B200 design (HBM-bound: 8 B per element):
//   * cols <= 128 (the detector's 80 classes): a warp owns a tile of 32 rows.
//     The tile is loaded flat and fully coalesced into shared memory with an odd
//     row stride (conflict-free), then LANE r walks ROW r: the running max, the
//     first-max index and the index-ordered sum are sequential per row exactly as
//     in the reference loop, yet 32 rows advance in lock-step.  The normalised
//     tile is written back flat and coalesced.
//   * larger rows: a warp per row; exponentials land in a per-warp shared buffer
//     and are summed in index order (in 4096-element chunks for longer rows).
//   * exp is tk_dm_expf (bit-identical to the oracle's deterministic mode and
//     within 2 ulp of libm, inside the reference test's 1e-6 absolute bound).
#include "tk_common.cuh"

namespace tkb {

static constexpr int kRowThreads = 256;
static constexpr int kRowWarps = kRowThreads / 32;
static constexpr uint32_t kTileMaxCols = 128;
static constexpr uint32_t kSeqMaxCols = 4096;

struct RowK {
    const float* in;
    float* out;          // may be null
    uint32_t* amax;      // may be null
    uint32_t rows, cols, stride;   // stride: odd smem row stride (tile kernel)
    uint32_t ntiles;
    float one;                     // 1.0f at run time (sum2 in tk_common.cuh)
};

// tk_dm_expf (include/tk_b200_detmath.h) for two arguments at once in packed fp32: every lane runs
// the scalar operation sequence (t = x*log2e; nf = (t + 1.5*2^23) - 1.5*2^23; r = x - nf*ln2_hi - nf*ln2_lo
// in two FMAs; degree-6 Horner; two exact power-of-two scalings), so the result is bit-identical to it.
__device__ __forceinline__ float scale_pow2(int n) { return __uint_as_float((uint32_t)(n + 127) << 23); }
__device__ __forceinline__ float2 dm_expf2(float2 x, float one) {
    const float2 t = mul2(x, f2s(1.44269502162933349609375f));
    const float2 nf = add2(sum2(t, f2s(12582912.0f), one), f2s(-12582912.0f));
    float2 r = fma2(nf, f2s(-0.693145751953125f), x);
    r = fma2(nf, f2s(-1.42860676533018704503775e-06f), r);
    float2 p = f2s(1.0f / 720.0f);
    p = fma2(p, r, f2s(1.0f / 120.0f));
    p = fma2(p, r, f2s(1.0f / 24.0f));
    p = fma2(p, r, f2s(1.0f / 6.0f));
    p = fma2(p, r, f2s(0.5f));
    p = fma2(p, r, f2s(1.0f));
    p = fma2(p, r, f2s(1.0f));
    const int nx = (int)nf.x, ny = (int)nf.y;
    const int nx1 = nx / 2, ny1 = ny / 2;
    float2 e = mul2(mul2(p, f2(scale_pow2(nx1), scale_pow2(ny1))), f2(scale_pow2(nx - nx1), scale_pow2(ny - ny1)));
    const float hi = 88.72283935546875f, lo = -103.97208404541015625f;
    e.x = (x.x != x.x) ? x.x : (x.x > hi ? __int_as_float(0x7f800000) : (x.x < lo ? 0.0f : e.x));
    e.y = (x.y != x.y) ? x.y : (x.y > hi ? __int_as_float(0x7f800000) : (x.y < lo ? 0.0f : e.y));
    return e;
}

// e / s, correctly rounded, from the row's correctly rounded reciprocal rs = RN(1/s): q = e*rs,
// r = e - q*s (exact in one FMA), q + r*rs (Markstein).  Exactness of r needs e well above the
// subnormal range; tiny e, non-finite s and NaNs take the IEEE divide.
__device__ __forceinline__ float div_by_row_sum(float e, float s, float rs) {
    if (e >= 1e-30f && s <= 3.0e38f) {
        const float q = __fmul_rn(e, rs);
        const float r = __fmaf_rn(-q, s, e);
        return __fmaf_rn(r, rs, q);
    }
    return __fdiv_rn(e, s);
}

// VEC: cols % 4 == 0 and both tensors 16-byte aligned — the flat passes move float4s (a float4 never
// straddles a row then); otherwise one element per step.
template <bool VEC>
__global__ void __launch_bounds__(kRowThreads) softmax_tile_kernel(const RowK k) {
    extern __shared__ __align__(16) float tsm[];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* tile = tsm + (size_t)warp * 32 * k.stride;
    constexpr uint32_t W = VEC ? 4u : 1u;                         // elements per lane per flat step
    for (uint32_t t = blockIdx.x * kRowWarps + warp; t < k.ntiles; t += gridDim.x * kRowWarps) {
        const uint32_t r0 = t * 32;
        const uint32_t nr = min(32u, k.rows - r0);
        const uint32_t nel = nr * k.cols;
        const float* src = k.in + (size_t)r0 * k.cols;
        // flat coalesced load; (row, col) advanced incrementally (32*W = a*cols + b)
        const uint32_t step_r = (32 * W) / k.cols, step_c = (32 * W) % k.cols;
        uint32_t rr = (lane * W) / k.cols, cc = (lane * W) % k.cols;
        for (uint32_t e = lane * W; e < nel; e += 32 * W) {
            float* dst = tile + rr * k.stride + cc;
            if constexpr (VEC) {
                const float4 v = __ldcs(reinterpret_cast<const float4*>(src + e));
                dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
            } else {
                dst[0] = __ldcs(src + e);
            }
            rr += step_r; cc += step_c;
            if (cc >= k.cols) { cc -= k.cols; ++rr; }
        }
        __syncwarp();
        float sum = 1.0f, rsum = 1.0f;
        if (lane < nr) {
            float* row = tile + lane * k.stride;
            float mx = row[0];
            uint32_t best = 0;
            for (uint32_t j = 1; j < k.cols; ++j) {
                float v = row[j];
                if (v > mx) { mx = v; best = j; }
            }
            if (k.amax) k.amax[r0 + lane] = best;
            if (k.out) {
                // four exponentials in flight per step (two packed pairs), the sum strictly in index order
                sum = 0.0f;
                const float2 nmx = f2s(-mx);                   // x - mx == x + (-mx)
                uint32_t j = 0;
                for (; j + 4 <= k.cols; j += 4) {
                    const float2 ea = dm_expf2(add2(f2(row[j], row[j + 1]), nmx), k.one);
                    const float2 eb = dm_expf2(add2(f2(row[j + 2], row[j + 3]), nmx), k.one);
                    row[j] = ea.x; row[j + 1] = ea.y; row[j + 2] = eb.x; row[j + 3] = eb.y;
                    sum = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(sum, ea.x), ea.y), eb.x), eb.y);
                }
                for (; j < k.cols; ++j) {
                    float e = tk_dm_expf(__fsub_rn(row[j], mx));
                    row[j] = e;
                    sum = __fadd_rn(sum, e);
                }
                rsum = __frcp_rn(sum);
            }
        }
        __syncwarp();
        if (k.out) {
            float* dst = k.out + (size_t)r0 * k.cols;
            rr = (lane * W) / k.cols; cc = (lane * W) % k.cols;
            for (uint32_t e0 = 0; e0 < nel; e0 += 32 * W) {      // uniform trip count: the shuffles need all lanes
                const uint32_t e = e0 + lane * W;
                const float s = __shfl_sync(0xffffffffu, sum, rr & 31);
                const float rs = __shfl_sync(0xffffffffu, rsum, rr & 31);
                if (e < nel) {
                    const float* sp = tile + rr * k.stride + cc;
                    if constexpr (VEC) {
                        __stcs(reinterpret_cast<float4*>(dst + e),
                               make_float4(div_by_row_sum(sp[0], s, rs), div_by_row_sum(sp[1], s, rs),
                                           div_by_row_sum(sp[2], s, rs), div_by_row_sum(sp[3], s, rs)));
                    } else {
                        __stcs(dst + e, div_by_row_sum(sp[0], s, rs));
                    }
                }
                rr += step_r; cc += step_c;
                if (cc >= k.cols) { cc -= k.cols; ++rr; }
            }
        }
        __syncwarp();
    }
}

__global__ void __launch_bounds__(kRowThreads) softmax_row_kernel(const RowK k) {
    extern __shared__ __align__(16) float rsm[];
    const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool seq = k.cols <= kSeqMaxCols;
    float* ebuf = rsm + (size_t)warp * (seq ? k.cols : kSeqMaxCols);
    for (uint32_t r = blockIdx.x * kRowWarps + warp; r < k.rows; r += gridDim.x * kRowWarps) {
        const float* row = k.in + (size_t)r * k.cols;
        // sequential-scan semantics: mx = x[0]; if (x[j] > mx) ...  (NaN at 0 sticks, NaN elsewhere ignored)
        float mv = 0.0f; uint32_t mi = 0xffffffffu;
        for (uint32_t j = lane; j < k.cols; j += 32) {
            float v = __ldg(row + j);
            if (j == 0 || (v == v && !(mv != mv) && (mi == 0xffffffffu || v > mv))) { mv = v; mi = j; }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            float ov = __shfl_xor_sync(0xffffffffu, mv, d);
            uint32_t oi = __shfl_xor_sync(0xffffffffu, mi, d);
            if (oi == 0xffffffffu || (mi != 0xffffffffu && mv != mv)) continue;
            if (mi == 0xffffffffu || ov != ov || ov > mv || (ov == mv && oi < mi)) { mv = ov; mi = oi; }
        }
        if (lane == 0 && k.amax) k.amax[r] = mi;
        if (!k.out) continue;
        float sum = 0.0f;
        if (seq) {
            for (uint32_t j = lane; j < k.cols; j += 32) ebuf[j] = tk_dm_expf(__fsub_rn(__ldg(row + j), mv));
            __syncwarp();
            if (lane == 0) for (uint32_t j = 0; j < k.cols; ++j) sum = __fadd_rn(sum, ebuf[j]);
            sum = __shfl_sync(0xffffffffu, sum, 0);
            for (uint32_t j = lane; j < k.cols; j += 32) __stcs(k.out + (size_t)r * k.cols + j, __fdiv_rn(ebuf[j], sum));
            __syncwarp();
        } else {
            // Long rows: the exponentials pass through the per-warp buffer in chunks so the
            // sum still runs strictly in index order (one running fp32 accumulator, as
            // tests/tk_gpu_softmax_test.cpp:41-44); the output pass recomputes them.
            for (uint32_t c0 = 0; c0 < k.cols; c0 += kSeqMaxCols) {
                const uint32_t n = min(kSeqMaxCols, k.cols - c0);
                for (uint32_t j = lane; j < n; j += 32) ebuf[j] = tk_dm_expf(__fsub_rn(__ldg(row + c0 + j), mv));
                __syncwarp();
                if (lane == 0) for (uint32_t j = 0; j < n; ++j) sum = __fadd_rn(sum, ebuf[j]);
                __syncwarp();
            }
            sum = __shfl_sync(0xffffffffu, sum, 0);
            for (uint32_t j = lane; j < k.cols; j += 32)
                __stcs(k.out + (size_t)r * k.cols + j, __fdiv_rn(tk_dm_expf(__fsub_rn(__ldg(row + j), mv)), sum));
        }
    }
}

static tk_error_code_t launch_rows(const float* in, float* out, uint32_t* amax, uint32_t rows, uint32_t cols, cudaStream_t stream) {
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    RowK k{};
    k.in = in; k.out = out; k.amax = amax; k.rows = rows; k.cols = cols; k.one = 1.0f;
    if (cols <= kTileMaxCols) {
        k.stride = cols | 1u;
        k.ntiles = (rows + 31) / 32;
        const size_t smem = (size_t)kRowWarps * 32 * k.stride * 4;
        const bool vec = (cols % 4 == 0) && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0;
        auto kern = vec ? softmax_tile_kernel<true> : softmax_tile_kernel<false>;
        TK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), TK_ERROR_GPU_CUDA_ERROR);
        int per_sm = 0;             // one wave of persistent CTAs: what registers AND the tile's shared memory let an SM hold
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kRowThreads, smem) != cudaSuccess || per_sm < 1) {
            cudaGetLastError();
            per_sm = 1;
        }
        uint32_t grid = (k.ntiles + kRowWarps - 1) / kRowWarps;
        const uint32_t cap = (uint32_t)dev->sm_count * (uint32_t)per_sm;
        if (grid > cap) grid = cap;
        kern<<<grid, kRowThreads, smem, stream>>>(k);
        return launch_status("softmax_tile_kernel");
    }
    const size_t smem = (size_t)kRowWarps * (cols <= kSeqMaxCols ? cols : kSeqMaxCols) * 4;
    if (smem > 48 * 1024)
        TK_CUDA_TRY(cudaFuncSetAttribute(softmax_row_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), TK_ERROR_GPU_CUDA_ERROR);
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, softmax_row_kernel, kRowThreads, smem) != cudaSuccess || per_sm < 1) {
        cudaGetLastError();
        per_sm = 1;
    }
    uint32_t grid = (rows + kRowWarps - 1) / kRowWarps;
    const uint32_t cap = (uint32_t)dev->sm_count * (uint32_t)per_sm;
    if (grid > cap) grid = cap;
    softmax_row_kernel<<<grid, kRowThreads, smem, stream>>>(k);
    return launch_status("softmax_row_kernel");
}

}  // namespace tkb

using namespace tkb;

// src/gpu/cuda/tk_cuda_kernels.cu:362-393 (argument checks and codes kept; any num_cols accepted).
extern "C" tk_error_code_t tk_kernels_softmax(const tk_softmax_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_input_tensor || !p->d_output_tensor) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_rows == 0 || p->num_cols == 0) return TK_ERROR_INVALID_DIMENSIONS;
    return launch_rows(static_cast<const float*>(p->d_input_tensor), static_cast<float*>(p->d_output_tensor), nullptr,
                       p->num_rows, p->num_cols, stream);
}

extern "C" tk_error_code_t tk_kernels_argmax_rows(const tk_b200_argmax_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_input || (!p->d_softmax && !p->d_argmax)) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_rows == 0 || p->num_cols == 0) return TK_ERROR_INVALID_DIMENSIONS;
    return launch_rows(p->d_input, p->d_softmax, p->d_argmax, p->num_rows, p->num_cols, stream);
}
