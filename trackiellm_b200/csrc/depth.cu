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
//       Persistent CTAs draw (frame, detection) items from a ticket counter; the samples of an item are
//       compacted into shared memory, sorted by a normalised bitonic network held in registers
//       (in-thread min/max pairs, warp shuffles, shared memory only for the widest strides; any
//       length, +inf padding), and summed in ascending order by one thread so the fp32 result is
//       bit-identical to the sequential reference.
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
        auto take = [&](uint32_t i, const float4 v) {
            if (i - lo < k.cache_floats) *reinterpret_cast<float4*>(cache + (i - lo)) = v;
            mn = fminf(fminf(mn, v.x), fminf(v.y, fminf(v.z, v.w)));
            mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
        };
        constexpr uint32_t S = 4 * kMetricThreads;
        uint32_t i = lo + 4 * tid;
        for (; i + 3 * S < hi; i += 4 * S) {                      // four 128-bit loads in flight per thread
            const float4 a = __ldcs(reinterpret_cast<const float4*>(in + i));
            const float4 b = __ldcs(reinterpret_cast<const float4*>(in + i + S));
            const float4 c = __ldcs(reinterpret_cast<const float4*>(in + i + 2 * S));
            const float4 d = __ldcs(reinterpret_cast<const float4*>(in + i + 3 * S));
            take(i, a); take(i + S, b); take(i + 2 * S, c); take(i + 3 * S, d);
        }
        for (; i < hi; i += S) take(i, __ldcs(reinterpret_cast<const float4*>(in + i)));
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

constexpr int kDistThreads = 512;
constexpr uint32_t kDistFrames = 1024;     // frames per launch (their count prefix lives in shared memory)
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
    uint32_t* ticket;       // work counter of this launch (zeroed before it)
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

// ---- register-resident sort of up to 512 * E samples ------------------------------------------
// Thread t owns the E consecutive elements t*E .. t*E+E-1 (+inf beyond n).  Normalised bitonic
// network: strides below E are compile-time min/max pairs in registers (2 instructions per
// compare-exchange), thread strides 1..16 are warp shuffles, 32..256 go through shared memory.
template <int E>
__device__ __forceinline__ void regs_flip(float (&v)[E], int kk) {          // blocks of kk <= E inside the thread
#pragma unroll
    for (int e = 0; e < E; ++e) {
        const int off = e & (kk - 1), pe = (e & ~(kk - 1)) + kk - 1 - off;
        if (e < pe) { const float lo = fminf(v[e], v[pe]), hi = fmaxf(v[e], v[pe]); v[e] = lo; v[pe] = hi; }
    }
}
template <int E>
__device__ __forceinline__ void regs_half(float (&v)[E], int j) {           // stride j < E inside the thread
#pragma unroll
    for (int e = 0; e < E; ++e) {
        if ((e & j) == 0) { const float lo = fminf(v[e], v[e | j]), hi = fmaxf(v[e], v[e | j]); v[e] = lo; v[e | j] = hi; }
    }
}
// Compare-exchange with thread tid ^ mask; `rev`: the partner's elements in reverse order (flip step).
template <int E>
__device__ __forceinline__ void regs_cross(float (&v)[E], uint32_t tid, uint32_t mask, bool rev, float* xbuf) {
    float o[E];
    if (mask < 32u) {
#pragma unroll
        for (int e = 0; e < E; ++e) {
            const float fwd = __shfl_xor_sync(0xffffffffu, v[e], mask);
            const float bwd = __shfl_xor_sync(0xffffffffu, v[E - 1 - e], mask);
            o[e] = rev ? bwd : fwd;
        }
    } else {
#pragma unroll
        for (int e = 0; e < E; ++e) xbuf[e * kDistThreads + tid] = v[e];
        __syncthreads();
#pragma unroll
        for (int e = 0; e < E; ++e) o[e] = xbuf[(rev ? E - 1 - e : e) * kDistThreads + (tid ^ mask)];
        __syncthreads();
    }
    const bool lower = tid < (tid ^ mask);
#pragma unroll
    for (int e = 0; e < E; ++e) v[e] = lower ? fminf(v[e], o[e]) : fmaxf(v[e], o[e]);
}
template <int E>
__device__ void sort_regs(float* a, uint32_t n, uint32_t tid) {
    float v[E];
#pragma unroll
    for (int e = 0; e < E; ++e) { const uint32_t g = tid * E + e; v[e] = g < n ? a[g] : __int_as_float(0x7f800000); }
    __syncthreads();                                           // `a` doubles as the exchange buffer
#pragma unroll
    for (int kk = 2; kk <= E; kk <<= 1) {
        regs_flip<E>(v, kk);
#pragma unroll
        for (int j = kk >> 2; j >= 1; j >>= 1) regs_half<E>(v, j);
    }
    const uint32_t tmax = (((1u << log2_ceil(n)) + E - 1) / E);   // threads holding real samples, a power of two
    for (uint32_t kt = 2; kt <= tmax; kt <<= 1) {
        regs_cross<E>(v, tid, kt - 1, true, a);
        for (uint32_t jt = kt >> 2; jt >= 1; jt >>= 1) regs_cross<E>(v, tid, jt, false, a);
#pragma unroll
        for (int j = E >> 1; j >= 1; j >>= 1) regs_half<E>(v, j);
    }
#pragma unroll
    for (int e = 0; e < E; ++e) { const uint32_t g = tid * E + e; if (g < n) a[g] = v[e]; }
    __syncthreads();
}
// Bitonic clean-up of a half-cleaned block of 512 * E slots: thread strides jt0 .. 1, then the in-register strides.
template <int E>
__device__ void clean_regs(float* a, uint32_t n, uint32_t tid, uint32_t jt0) {
    float v[E];
#pragma unroll
    for (int e = 0; e < E; ++e) { const uint32_t g = tid * E + e; v[e] = g < n ? a[g] : __int_as_float(0x7f800000); }
    __syncthreads();
    for (uint32_t jt = jt0; jt >= 1; jt >>= 1) regs_cross<E>(v, tid, jt, false, a);
#pragma unroll
    for (int j = E >> 1; j >= 1; j >>= 1) regs_half<E>(v, j);
#pragma unroll
    for (int e = 0; e < E; ++e) { const uint32_t g = tid * E + e; if (g < n) a[g] = v[e]; }
    __syncthreads();
}
__device__ void sort_small(float* a, uint32_t n, uint32_t tid) {           // n <= 16384, a in shared memory (>= 512 floats)
    const uint32_t np2 = 1u << log2_ceil(n);
    if (np2 <= 512) sort_regs<1>(a, n, tid);
    else if (np2 <= 1024) sort_regs<2>(a, n, tid);
    else if (np2 <= 2048) sort_regs<4>(a, n, tid);
    else if (np2 <= 4096) sort_regs<8>(a, n, tid);
    else if (np2 <= 8192) sort_regs<16>(a, n, tid);
    else {
        // 8192 < n <= 16384: both halves in registers, then the last merge — its flip and its
        // stride-4096 step in shared memory, the rest (strides 2048 .. 1) per half in registers.
        sort_regs<16>(a, 8192, tid);
        sort_regs<16>(a + 8192, n - 8192, tid);
        flip_step(a, n, 14, 8192, tid);
        __syncthreads();
        half_step(a, n, 12, 8192, tid);
        __syncthreads();
        clean_regs<16>(a, 8192, tid, 128);
        clean_regs<16>(a + 8192, n - 8192, tid, 128);
    }
}

// Ascending-order fp32 sum of s[0..m) by ONE thread (the rounding sequence of the reference's
// sequential sum); 128-bit shared-memory loads keep the dependent FADD chain fed.
__device__ __forceinline__ float seq_sum_shared(const float* s, uint32_t i0, uint32_t i1, float acc) {
    uint32_t i = i0;
    for (; i < i1 && (i & 3u); ++i) acc = __fadd_rn(acc, s[i]);
    for (; i + 4 <= i1; i += 4) {
        const float4 v = *reinterpret_cast<const float4*>(s + i);
        acc = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, v.x), v.y), v.z), v.w);
    }
    for (; i < i1; ++i) acc = __fadd_rn(acc, s[i]);
    return acc;
}

// One detection: the whole CTA works on (frame, slot).  Every branch below is CTA-uniform.
__device__ void distance_item(const DistK& k, uint32_t frame, uint32_t slot, float* sbuf) {
    __shared__ uint32_t s_cnt, s_lo, s_hi, s_slot;
    __shared__ float s_acc;
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    float* outp = k.out + (size_t)frame * k.max_det + slot;

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
            uint32_t s = (frame + slot) % kDistSlots;
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
        sort_small(a, n, tid);
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
            const float acc = seq_sum_shared(sbuf, i0, i1, 0.0f);
            *outp = i1 > i0 ? __fdiv_rn(acc, __uint2float_rn(i1 - i0)) : -1.0f;
        }
    } else {
        if (tid == 0) s_acc = 0.0f;
        for (uint32_t c0 = i0; c0 < i1; c0 += k.smem_floats) {
            const uint32_t m = min(k.smem_floats, i1 - c0);
            __syncthreads();
            for (uint32_t i = tid; i < m; i += kDistThreads) sbuf[i] = a[c0 + i];
            __syncthreads();
            if (tid == 0) s_acc = seq_sum_shared(sbuf, 0, m, s_acc);
        }
        __syncthreads();
        if (tid == 0) {
            *outp = i1 > i0 ? __fdiv_rn(s_acc, __uint2float_rn(i1 - i0)) : -1.0f;
            __threadfence();
            atomicExch(&k.locks[s_slot], 0u);
        }
    }
}

// Persistent CTAs over the flattened list of real detections of frames [frame0, frame0 + nf):
// the per-frame counts are prefix-summed in shared memory, items are handed out dynamically,
// and the unused slots of the output are filled with -1 on the side.
__global__ void __launch_bounds__(kDistThreads, 2) detection_distance_kernel(const DistK k, uint32_t frame0, uint32_t nf) {
    extern __shared__ __align__(16) float sbuf[];
    __shared__ uint32_t pre[kDistFrames + 1];
    const uint32_t tid = threadIdx.x;
    for (uint32_t f = tid; f < nf; f += kDistThreads)
        pre[f + 1] = min(k.counts ? k.counts[frame0 + f] : k.max_det, k.max_det);
    if (tid == 0) pre[0] = 0;
    __syncthreads();
    if (tid < 32) {                                    // inclusive scan of pre[1..nf] by one warp, 32 at a time
        uint32_t carry = 0;
        for (uint32_t b = 0; b < nf; b += 32) {
            const uint32_t f = b + tid;
            uint32_t v = f < nf ? pre[f + 1] : 0u;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
                if ((int)tid >= o) v += u;
            }
            if (f < nf) pre[f + 1] = v + carry;
            carry += __shfl_sync(0xffffffffu, v, 31);
        }
    }
    __syncthreads();
    // unused slots -> -1
    for (size_t i = (size_t)blockIdx.x * kDistThreads + tid; i < (size_t)nf * k.max_det; i += (size_t)gridDim.x * kDistThreads) {
        const uint32_t f = (uint32_t)(i / k.max_det), sl = (uint32_t)(i % k.max_det);
        if (sl >= pre[f + 1] - pre[f]) k.out[(size_t)(frame0 + f) * k.max_det + sl] = -1.0f;
    }
    // Items cost anything from a few hundred to 16 K samples: CTAs draw them from a ticket counter.
    __shared__ uint32_t s_item;
    const uint32_t total = pre[nf];
    for (;;) {
        __syncthreads();                               // the previous item's shared state is dead
        if (tid == 0) s_item = atomicAdd(k.ticket, 1u);
        __syncthreads();
        const uint32_t item = s_item;
        if (item >= total) break;
        uint32_t lo = 0, hi = nf;                      // largest f with pre[f] <= item
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (pre[mid] <= item) lo = mid; else hi = mid; }
        distance_item(k, frame0 + lo, item - pre[lo], sbuf);
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
    if (n > 0x7fffffffull) return TK_ERROR_INVALID_DIMENSIONS;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    DistK k{};
    k.dets = p->d_detections; k.counts = p->d_num_detections; k.depth = p->d_depth_maps; k.out = p->d_distances;
    k.num_frames = p->num_frames; k.max_det = p->max_detections;
    k.depth_w = p->depth_width; k.depth_h = p->depth_height; k.frame_w = p->frame_width; k.frame_h = p->frame_height;
    uint32_t cap = 512;                                    // >= one exchange slot per thread
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
    uint32_t per_sm = (uint32_t)((size_t)200 * 1024 / (smem + 8 * 1024));
    per_sm = per_sm < 1 ? 1 : (per_sm > 2048 / kDistThreads ? 2048 / kDistThreads : per_sm);
    const uint32_t launches = (p->num_frames + kDistFrames - 1) / kDistFrames;
    uint32_t* tickets = nullptr;                           // one counter per launch, stream-ordered lifetime
    TK_CUDA_TRY(cudaMallocAsync(reinterpret_cast<void**>(&tickets), 4 * (size_t)launches, stream), TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED);
    TK_CUDA_TRY(cudaMemsetAsync(tickets, 0, 4 * (size_t)launches, stream), TK_ERROR_GPU_CUDA_ERROR);
    tk_error_code_t rc = TK_SUCCESS;
    for (uint32_t f0 = 0; f0 < p->num_frames && rc == TK_SUCCESS; f0 += kDistFrames) {
        k.ticket = tickets + f0 / kDistFrames;
        const uint32_t nf = p->num_frames - f0 < kDistFrames ? p->num_frames - f0 : kDistFrames;
        unsigned long long items = (unsigned long long)nf * p->max_detections;
        const unsigned long long cap = (unsigned long long)dev->sm_count * per_sm;
        detection_distance_kernel<<<(unsigned)(items < cap ? items : cap), kDistThreads, smem, stream>>>(k, f0, nf);
        rc = launch_status("detection_distance_kernel");
    }
    cudaFreeAsync(tickets, stream);
    return rc;
}
