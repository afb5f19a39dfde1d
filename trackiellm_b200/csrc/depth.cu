// depth.cu — depth-estimator post-processing either side of the detector path
// (additive entry points, same conventions).
//
//   tk_kernels_inverse_depth_to_metric : MiDaS inverse depth -> metric depth,
//       convert_inverse_depth_to_metric of src/vision/tk_depth_midas.c:473-499
//       (min/max scan, then max_depth - (d - min)/(max - min) * (max_depth - min_depth)).
//       One thread-block cluster of 8 CTAs per map: every CTA keeps its slice of the
//       map in shared memory while the 8 partial extrema are exchanged through
//       distributed shared memory, so the map is read from HBM once and written once
//       (the reference walks it three times).
//   tk_kernels_detection_distance : per-detection robust distance,
//       calculate_raw_distance of src/vision/src/object_analysis.rs:227-279
//       (bbox -> depth window -> valid samples -> sort -> IQR filter -> mean).
//       One CTA per detection; samples sorted in shared memory by a normalised
//       bitonic network (any length, no padding), summed in ascending order by one
//       thread so the fp32 result is bit-identical to the sequential reference.
#include "tk_common.cuh"

#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace tkb {

// ------------------------------------------------------------------------- //
// inverse depth -> metric                                                     //
// ------------------------------------------------------------------------- //
constexpr int kMetricThreads = 512;
constexpr int kMetricCluster = 8;

struct MetricK {
    const float* in;
    float* out;
    uint32_t n;             // elements per map
    uint32_t num_maps;
    uint32_t slice;         // elements per CTA (multiple of 4)
    uint32_t cache_floats;  // elements of the slice kept in shared memory (multiple of 4)
    float min_depth, max_depth;
    int vec;                // 1: n % 4 == 0 and both bases 16-byte aligned
};

__device__ __forceinline__ float metric_of(float d, float mn, float range, float max_depth, float span, bool flat) {
    if (flat) return max_depth;                                   // tk_depth_midas.c:485-491
    const float normalized = __fdiv_rn(__fsub_rn(d, mn), range);  // :496
    return __fsub_rn(max_depth, __fmul_rn(normalized, span));     // :497
}

__global__ void __launch_bounds__(kMetricThreads) inverse_depth_metric_kernel(const MetricK k) {
    extern __shared__ __align__(16) float cache[];
    __shared__ float wmin[kMetricThreads / 32], wmax[kMetricThreads / 32];
    __shared__ float cmin[kMetricCluster], cmax[kMetricCluster];   // written by every CTA of the cluster

    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t rank = cluster.block_rank();
    const uint32_t map = blockIdx.x / kMetricCluster;
    const float* in = k.in + (size_t)map * k.n;
    float* out = k.out + (size_t)map * k.n;
    const uint32_t lo = min(rank * k.slice, k.n), hi = min(lo + k.slice, k.n);
    const uint32_t tid = threadIdx.x;

    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");   // resident (waited on before DSMEM)

    // pass 1: stream the slice in, keep it in shared memory, scan for the extrema.
    // `if (d < min) min = d` never takes a NaN (tk_depth_midas.c:479-482); fminf/fmaxf drop NaNs likewise.
    float mn = __int_as_float(0x7f800000), mx = __int_as_float(0xff800000);
    if (k.vec) {
        for (uint32_t i = lo + 4 * tid; i < hi; i += 4 * kMetricThreads) {
            const float4 v = __ldcs(reinterpret_cast<const float4*>(in + i));
            if (i - lo < k.cache_floats) *reinterpret_cast<float4*>(cache + (i - lo)) = v;
            mn = fminf(fminf(mn, v.x), fminf(v.y, fminf(v.z, v.w)));
            mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
        }
    } else {
        for (uint32_t i = lo + tid; i < hi; i += kMetricThreads) {
            const float v = in[i];
            if (i - lo < k.cache_floats) cache[i - lo] = v;
            mn = fminf(mn, v); mx = fmaxf(mx, v);
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if ((tid & 31) == 0) { wmin[tid >> 5] = mn; wmax[tid >> 5] = mx; }
    __syncthreads();
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");             // every peer has started
    if (tid < 32) {
        mn = tid < kMetricThreads / 32 ? wmin[tid] : __int_as_float(0x7f800000);
        mx = tid < kMetricThreads / 32 ? wmax[tid] : __int_as_float(0xff800000);
#pragma unroll
        for (int o = 8; o; o >>= 1) {
            mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        // The reference seeds both extrema with depth_data[0] (:476-477): a NaN there sticks.
        // Element 0 sits in rank 0's cache; its partial carries the NaN to every peer.
        if (rank == 0 && cache[0] != cache[0]) { mn = cache[0]; mx = cache[0]; }
        if (tid < kMetricCluster) {                                         // DSMEM store into peer `tid`
            *cluster.map_shared_rank(&cmin[rank], tid) = mn;
            *cluster.map_shared_rank(&cmax[rank], tid) = mx;
        }
    }
    cluster.sync();
    mn = cmin[0]; mx = cmax[0];
    if (mn == mn) {
#pragma unroll
        for (int r = 1; r < kMetricCluster; ++r) { mn = fminf(mn, cmin[r]); mx = fmaxf(mx, cmax[r]); }
    }

    const float range = __fsub_rn(mx, mn);
    const bool flat = range < 1e-6f;                                        // :485 (false for NaN)
    const float span = __fsub_rn(k.max_depth, k.min_depth);

    // pass 2: from shared memory (or, beyond the cache, from L2) to HBM.
    if (k.vec) {
        for (uint32_t i = lo + 4 * tid; i < hi; i += 4 * kMetricThreads) {
            float4 v = (i - lo < k.cache_floats) ? *reinterpret_cast<const float4*>(cache + (i - lo))
                                                 : __ldcs(reinterpret_cast<const float4*>(in + i));
            v.x = metric_of(v.x, mn, range, k.max_depth, span, flat);
            v.y = metric_of(v.y, mn, range, k.max_depth, span, flat);
            v.z = metric_of(v.z, mn, range, k.max_depth, span, flat);
            v.w = metric_of(v.w, mn, range, k.max_depth, span, flat);
            __stcs(reinterpret_cast<float4*>(out + i), v);
        }
    } else {
        for (uint32_t i = lo + tid; i < hi; i += kMetricThreads) {
            const float v = (i - lo < k.cache_floats) ? cache[i - lo] : in[i];
            out[i] = metric_of(v, mn, range, k.max_depth, span, flat);
        }
    }
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_kernels_inverse_depth_to_metric(const tk_b200_depth_metric_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_inverse_depth || !p->d_metric_depth) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_maps == 0) return TK_SUCCESS;
    if (p->width == 0 || p->height == 0) return TK_ERROR_INVALID_DIMENSIONS;
    const unsigned long long n64 = (unsigned long long)p->width * p->height;
    if (n64 > 0x7fffffffull) return TK_ERROR_INVALID_DIMENSIONS;
    // In place is fine only when the slice cache holds everything; keep it simple: any overlap but identity is undefined.
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    MetricK k{};
    k.in = p->d_inverse_depth; k.out = p->d_metric_depth;
    k.n = (uint32_t)n64; k.num_maps = p->num_maps;
    k.slice = ((k.n + kMetricCluster - 1) / kMetricCluster + 3) & ~3u;
    const uint32_t cap = (96 * 1024) / 4;                                   // two CTAs per SM
    k.cache_floats = k.slice < cap ? k.slice : cap;
    k.min_depth = p->min_depth; k.max_depth = p->max_depth;
    k.vec = (k.n % 4 == 0) && ((reinterpret_cast<uintptr_t>(p->d_inverse_depth) | reinterpret_cast<uintptr_t>(p->d_metric_depth)) & 15) == 0;
    // in place beyond the cache would re-read converted values in pass 2
    if (p->d_inverse_depth == p->d_metric_depth && k.cache_floats < k.slice) {
        tk_error_set_detail("in-place conversion needs maps of at most %u elements", cap * kMetricCluster);
        return TK_ERROR_NOT_IMPLEMENTED;
    }
    const size_t smem = (size_t)k.cache_floats * 4;
    TK_CUDA_TRY(cudaFuncSetAttribute(inverse_depth_metric_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                TK_ERROR_GPU_CUDA_ERROR);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(kMetricCluster * k.num_maps);
    cfg.blockDim = dim3(kMetricThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = kMetricCluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    TK_CUDA_TRY(cudaLaunchKernelEx(&cfg, inverse_depth_metric_kernel, k), TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE);
    return launch_status("inverse_depth_metric_kernel");
}

// ------------------------------------------------------------------------- //
// per-detection distance (src/vision/src/object_analysis.rs:227-279)          //
// ------------------------------------------------------------------------- //
namespace tkb {

constexpr int kDistThreads = 256;
constexpr uint32_t kDistChunk = 16384;     // samples sorted in shared memory at once (64 KB)
constexpr uint32_t kDistSlots = 16;        // global scratch slots for windows with more samples
constexpr uint32_t kDistMinPoints = 10;    // MIN_DEPTH_POINTS_FOR_STATS, object_analysis.rs:233

struct DistK {
    const tk_b200_candidate_t* dets;
    const uint32_t* counts;
    const float* depth;
    float* out;
    uint32_t* locks;        // kDistSlots words, zeroed before the launch (or NULL)
    float* scratch;         // kDistSlots x depth_w*depth_h floats (or NULL)
    uint32_t num_frames, max_det, depth_w, depth_h, frame_w, frame_h;
    uint32_t smem_floats;   // capacity of the shared-memory sample buffer
};

__device__ __forceinline__ void cmp_swap(float* a, uint32_t lo, uint32_t hi) {
    const float x = a[lo], y = a[hi];
    if (y < x) { a[lo] = y; a[hi] = x; }
}
// Normalised bitonic network: every compare-exchange puts the smaller value at the
// lower index, so virtual +inf padding at indices >= n never moves and pairs that
// touch it are skipped — any n sorts without padding.
// flip step of merge size k = 1 << lk: i -> (blk*k + off, blk*k + k-1-off)
__device__ __forceinline__ void flip_step(float* a, uint32_t n, uint32_t lk, uint32_t half_np2, uint32_t tid) {
    const uint32_t hk = 1u << (lk - 1);
    for (uint32_t i = tid; i < half_np2; i += kDistThreads) {
        const uint32_t blk = i >> (lk - 1), off = i & (hk - 1);
        const uint32_t lo = (blk << lk) + off, hi = (blk << lk) + (hk << 1) - 1 - off;
        if (hi < n) cmp_swap(a, lo, hi);
    }
}
// half-cleaner with stride j = 1 << lj
__device__ __forceinline__ void half_step(float* a, uint32_t n, uint32_t lj, uint32_t half_np2, uint32_t tid) {
    const uint32_t j = 1u << lj;
    for (uint32_t i = tid; i < half_np2; i += kDistThreads) {
        const uint32_t lo = ((i & ~(j - 1)) << 1) | (i & (j - 1)), hi = lo + j;
        if (hi < n) cmp_swap(a, lo, hi);
    }
}
__device__ __forceinline__ uint32_t log2_ceil(uint32_t n) { return n <= 1 ? 0 : 32 - __clz(n - 1); }

// All merges k = 2 .. 2^lmax of `m` samples at `s` (shared memory).
__device__ void sort_shared(float* s, uint32_t m, uint32_t lmax, uint32_t tid) {
    const uint32_t half = (1u << log2_ceil(m)) >> 1;
    for (uint32_t lk = 1; lk <= lmax; ++lk) {
        flip_step(s, m, lk, half, tid);
        __syncthreads();
        for (int lj = (int)lk - 2; lj >= 0; --lj) { half_step(s, m, (uint32_t)lj, half, tid); __syncthreads(); }
    }
}

__global__ void __launch_bounds__(kDistThreads) detection_distance_kernel(const DistK k) {
    extern __shared__ __align__(16) float sbuf[];
    __shared__ uint32_t s_cnt, s_lo, s_hi, s_slot;
    __shared__ float s_acc;
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const uint32_t frame = blockIdx.y, slot = blockIdx.x;
    float* outp = k.out + (size_t)frame * k.max_det + slot;
    const uint32_t count = min(k.counts ? k.counts[frame] : k.max_det, k.max_det);
    if (slot >= count) { if (tid == 0) *outp = -1.0f; return; }

    // bbox -> depth-map window, object_analysis.rs:238-253 (f32 divide, multiply, round half
    // away from zero, saturating `as u32`; i32 additions wrap as in a release build)
    const tk_rect_t bb = k.dets[(size_t)frame * k.max_det + slot].bbox;
    const float fw = __uint2float_rn(k.frame_w), fh = __uint2float_rn(k.frame_h);
    const float wm1 = __uint2float_rn(k.depth_w - 1), hm1 = __uint2float_rn(k.depth_h - 1);
    const int xr = (int)((uint32_t)bb.x + (uint32_t)bb.w), yb = (int)((uint32_t)bb.y + (uint32_t)bb.h);
    const uint32_t x0 = __float2uint_rz(roundf(__fmul_rn(__fdiv_rn(__int2float_rn(bb.x), fw), wm1)));
    const uint32_t y0 = __float2uint_rz(roundf(__fmul_rn(__fdiv_rn(__int2float_rn(bb.y), fh), hm1)));
    const uint32_t x1 = __float2uint_rz(roundf(__fmul_rn(__fdiv_rn(__int2float_rn(xr), fw), wm1)));
    const uint32_t y1 = __float2uint_rz(roundf(__fmul_rn(__fdiv_rn(__int2float_rn(yb), fh), hm1)));
    if (x0 >= k.depth_w || y0 >= k.depth_h || x1 >= k.depth_w || y1 >= k.depth_h || x0 >= x1 || y0 >= y1) {
        if (tid == 0) *outp = -1.0f;
        return;
    }
    const float* dm = k.depth + (size_t)frame * k.depth_w * k.depth_h;
    const uint32_t wc = x1 - x0 + 1, total = wc * (y1 - y0 + 1);

    // pass 1: how many samples are valid (0.1 < d < 100, :260)
    if (tid == 0) { s_cnt = 0; s_lo = 0; s_hi = 0; }
    __syncthreads();
    uint32_t mine = 0;
    for (uint32_t t = tid; t < total; t += kDistThreads) {
        const float d = dm[(size_t)(y0 + t / wc) * k.depth_w + x0 + t % wc];
        mine += (d > 0.1f && d < 100.0f) ? 1u : 0u;
    }
    mine = __reduce_add_sync(0xffffffffu, mine);
    if (lane == 0 && mine) atomicAdd(&s_cnt, mine);
    __syncthreads();
    const uint32_t n = s_cnt;
    if (n < kDistMinPoints) { if (tid == 0) *outp = -1.0f; return; }
    __syncthreads();

    const bool big = n > k.smem_floats;
    float* a = sbuf;
    if (big) {
        if (!k.scratch) { if (tid == 0) *outp = __int_as_float(0x7fc00000); return; }   // launcher guarantees scratch
        if (tid == 0) {
            uint32_t s = (blockIdx.x + blockIdx.y) % kDistSlots;
            while (atomicCAS(&k.locks[s], 0u, 1u) != 0u) { s = (s + 1) % kDistSlots; __nanosleep(200); }
            __threadfence();
            s_slot = s;
        }
        __syncthreads();
        a = k.scratch + (size_t)s_slot * k.depth_w * k.depth_h;
    }

    // pass 2: gather the valid samples (any order — they are sorted next)
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    for (uint32_t t0 = 0; t0 < total; t0 += kDistThreads) {
        const uint32_t t = t0 + tid;
        float d = 0.0f;
        if (t < total) d = dm[(size_t)(y0 + t / wc) * k.depth_w + x0 + t % wc];
        const bool ok = d > 0.1f && d < 100.0f;
        const uint32_t m = __ballot_sync(0xffffffffu, ok);
        uint32_t base = 0;
        if (lane == 0 && m) base = atomicAdd(&s_cnt, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (ok) a[base + __popc(m & ((1u << lane) - 1))] = d;
    }
    __syncthreads();

    // sort ascending (:265)
    const uint32_t ln = log2_ceil(n);
    if (!big) {
        sort_shared(a, n, ln, tid);
    } else {
        const uint32_t lc = log2_ceil(k.smem_floats);           // smem_floats is a power of two here
        const uint32_t ch = 1u << lc, nch = (n + ch - 1) >> lc;
        const uint32_t half = (1u << ln) >> 1;
        // chunk-local merges k <= ch in shared memory
        for (uint32_t c = 0; c < nch; ++c) {
            const uint32_t m = min(ch, n - (c << lc));
            for (uint32_t i = tid; i < m; i += kDistThreads) sbuf[i] = a[(c << lc) + i];
            __syncthreads();
            sort_shared(sbuf, m, min(lc, log2_ceil(m)), tid);
            // a short last chunk still has to be a sorted run of the network at size ch: it is (sorted ascending)
            for (uint32_t i = tid; i < m; i += kDistThreads) a[(c << lc) + i] = sbuf[i];
            __syncthreads();
        }
        // merges across chunks: wide strides in global memory (L2), the rest per chunk in shared memory
        for (uint32_t lk = lc + 1; lk <= ln; ++lk) {
            flip_step(a, n, lk, half, tid);
            __syncthreads();
            int lj = (int)lk - 2;
            for (; lj >= (int)lc; --lj) { half_step(a, n, (uint32_t)lj, half, tid); __syncthreads(); }
            for (uint32_t c = 0; c < nch; ++c) {
                const uint32_t m = min(ch, n - (c << lc));
                for (uint32_t i = tid; i < m; i += kDistThreads) sbuf[i] = a[(c << lc) + i];
                __syncthreads();
                const uint32_t hm = (1u << log2_ceil(m)) >> 1;
                for (int l2 = lj; l2 >= 0; --l2) { half_step(sbuf, m, (uint32_t)l2, hm, tid); __syncthreads(); }
                for (uint32_t i = tid; i < m; i += kDistThreads) a[(c << lc) + i] = sbuf[i];
                __syncthreads();
            }
        }
    }

    // quartiles, IQR fence (:266-270); the filtered samples are a contiguous run [i0, i1)
    const float q1 = a[n / 4], q3 = a[n * 3 / 4];
    const float iqr = __fsub_rn(q3, q1);
    const float lb = __fsub_rn(q1, __fmul_rn(1.5f, iqr)), ub = __fadd_rn(q3, __fmul_rn(1.5f, iqr));
    uint32_t below = 0, upto = 0;
    for (uint32_t i = tid; i < n; i += kDistThreads) {
        const float d = a[i];
        below += d < lb ? 1u : 0u;
        upto += d <= ub ? 1u : 0u;
    }
    below = __reduce_add_sync(0xffffffffu, below);
    upto = __reduce_add_sync(0xffffffffu, upto);
    if (lane == 0) { if (below) atomicAdd(&s_lo, below); if (upto) atomicAdd(&s_hi, upto); }
    __syncthreads();
    const uint32_t i0 = s_lo, i1 = s_hi;

    // mean in ascending order, one thread: the fp32 rounding sequence of `iter().sum::<f32>()` (:273-275)
    if (!big) {
        if (tid == 0) {
            float acc = 0.0f;
            for (uint32_t i = i0; i < i1; ++i) acc = __fadd_rn(acc, a[i]);
            *outp = i1 > i0 ? __fdiv_rn(acc, __uint2float_rn(i1 - i0)) : -1.0f;
        }
    } else {
        if (tid == 0) s_acc = 0.0f;
        for (uint32_t c0 = i0; c0 < i1; c0 += k.smem_floats) {
            const uint32_t m = min(k.smem_floats, i1 - c0);
            __syncthreads();
            for (uint32_t i = tid; i < m; i += kDistThreads) sbuf[i] = a[c0 + i];
            __syncthreads();
            if (tid == 0) {
                float acc = s_acc;
                for (uint32_t i = 0; i < m; ++i) acc = __fadd_rn(acc, sbuf[i]);
                s_acc = acc;
            }
        }
        __syncthreads();
        if (tid == 0) {
            *outp = i1 > i0 ? __fdiv_rn(s_acc, __uint2float_rn(i1 - i0)) : -1.0f;
            __threadfence();
            atomicExch(&k.locks[s_slot], 0u);
        }
    }
}

}  // namespace tkb

extern "C" size_t tk_b200_detection_distance_workspace_bytes(uint32_t depth_width, uint32_t depth_height) {
    const unsigned long long n = (unsigned long long)depth_width * depth_height;
    if (n <= tkb::kDistChunk) return 0;
    return 256 + (size_t)tkb::kDistSlots * n * sizeof(float);
}

extern "C" tk_error_code_t tk_kernels_detection_distance(const tk_b200_distance_params_t* p, cudaStream_t stream) {
    using namespace tkb;
    if (!p || !p->d_detections || !p->d_depth_maps || !p->d_distances) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_frames == 0 || p->max_detections == 0) return TK_SUCCESS;
    if (p->depth_width == 0 || p->depth_height == 0 || p->frame_width == 0 || p->frame_height == 0)
        return TK_ERROR_INVALID_DIMENSIONS;
    const unsigned long long n = (unsigned long long)p->depth_width * p->depth_height;
    if (n > 0x7fffffffull || p->num_frames > 65535) return TK_ERROR_INVALID_DIMENSIONS;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    DistK k{};
    k.dets = p->d_detections; k.counts = p->d_num_detections; k.depth = p->d_depth_maps; k.out = p->d_distances;
    k.num_frames = p->num_frames; k.max_det = p->max_detections;
    k.depth_w = p->depth_width; k.depth_h = p->depth_height; k.frame_w = p->frame_width; k.frame_h = p->frame_height;
    uint32_t cap = 64;
    while (cap < n && cap < kDistChunk) cap <<= 1;
    k.smem_floats = cap;
    if (n > kDistChunk) {
        const size_t need = tk_b200_detection_distance_workspace_bytes(p->depth_width, p->depth_height);
        if (!p->d_workspace || p->workspace_bytes < need || (reinterpret_cast<uintptr_t>(p->d_workspace) & 15)) {
            tk_error_set_detail("depth maps of %llu samples need a %zu-byte workspace (tk_b200_detection_distance_workspace_bytes)", n, need);
            return TK_ERROR_BUFFER_TOO_SMALL;
        }
        k.locks = static_cast<uint32_t*>(p->d_workspace);
        k.scratch = reinterpret_cast<float*>(static_cast<uint8_t*>(p->d_workspace) + 256);
        TK_CUDA_TRY(cudaMemsetAsync(k.locks, 0, 256, stream), TK_ERROR_GPU_CUDA_ERROR);
    }
    const size_t smem = (size_t)cap * 4;
    TK_CUDA_TRY(cudaFuncSetAttribute(detection_distance_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                TK_ERROR_GPU_CUDA_ERROR);
    detection_distance_kernel<<<dim3(p->max_detections, p->num_frames), kDistThreads, smem, stream>>>(k);
    return launch_status("detection_distance_kernel");
}
