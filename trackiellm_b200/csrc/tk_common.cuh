// tk_common.cuh — shared host/device helpers for libtk_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/tk_b200_vision.h"
#include "../../include/tk_b200_detmath.h"

#define TK_CUDA_TRY(call, code)                                                     \
    do {                                                                            \
        cudaError_t e__ = (call);                                                   \
        if (e__ != cudaSuccess) {                                                   \
            tk_error_set_detail("%s:%d %s -> %s", __FILE__, __LINE__, #call,        \
                                cudaGetErrorString(e__));                           \
            return (code);                                                          \
        }                                                                           \
    } while (0)

namespace tkb {

struct DeviceInfo {
    int sm_count;
    int max_smem_optin;
    int cc_major;
};
// Cached per-device properties; returns nullptr when no usable CUDA device.
const DeviceInfo* device_info(int* out_device = nullptr);

// preprocess.cu: tk_b200_preprocess_batch; pdl_after_nms = launch as a programmatic dependent (only the pipeline's
// process_device passes true: there the kernel that precedes it on the stream is either its own nms_kernel or
// something that completes before this grid starts).  scan + d_scan_bitmap: run decode's pass 1 in an extra warp of
// every CTA (preprocess.cu, ScanK) and leave the ballot words in d_scan_bitmap [num_frames][ceil(P / 32)] for
// decode_detections_bitmap.
tk_error_code_t preprocess_batch_impl(const tk_b200_preprocess_batch_params_t* prm, cudaStream_t stream, bool pdl_after_nms,
                                      const tk_b200_decode_params_t* scan, uint32_t* d_scan_bitmap);
// preprocess.cu: see the definition.
bool preprocess_reads_rows_sparsely(const tk_b200_preprocess_batch_params_t* prm);

// decode.cu: tk_kernels_decode_detections with the objectness column supplied as a dense device array [frames][P]
// (nullptr: read from the rows).  The pipeline's host-logits path packs that column on the host so the rows can stay in
// page-locked host memory: pass 1 reads HBM, only the surviving rows cross PCIe.
tk_error_code_t decode_detections_packed(const tk_b200_decode_params_t* p, const float* d_objectness, cudaStream_t stream);

// decode.cu: decode_kernel with the NMS tail (layout 0 of tk_kernels_postprocess_fused); nms_k points at a filled NmsK.
tk_error_code_t decode_nms_cluster(const tk_b200_decode_params_t* p, const float* d_objectness, const void* nms_k,
                                   void* scratch, uint32_t scratch_pitch, cudaStream_t stream);

// post.cu: tk_kernels_postprocess_fused with the objectness column supplied densely (see decode_detections_packed).
tk_error_code_t postprocess_fused_packed(const tk_b200_post_params_t* p, const float* d_objectness, cudaStream_t stream);

// post.cu: decode + NMS of one frame with no candidate cap (see the definition).
tk_error_code_t postprocess_frame_uncapped(const float* d_logits, uint32_t P, uint32_t C, tk_b200_activation_e act,
                                           float conf_thr, float iou_thr, uint32_t max_det, int class_agnostic,
                                           const uint32_t restore[4], tk_b200_candidate_t* d_dets, uint32_t* d_ndets,
                                           cudaStream_t st);

// decode.cu: decode_kernel with pass 1 already done — d_bitmap [frames][ceil(P / 32)] as the pre-process kernel's
// scanner warps leave it.
tk_error_code_t decode_detections_bitmap(const tk_b200_decode_params_t* p, const uint32_t* d_bitmap, cudaStream_t stream);

// Programmatic-dependent-launch chain of one pipeline step (pre-process -> decode -> NMS on one stream).  The pipeline's
// process_device sets this thread-local hint around its three launches: the pre-process kernel then announces its
// dependents at its top, and decode_kernel / nms_kernel are launched with the programmatic-serialization attribute.  Both
// start with griddepcontrol.wait as their FIRST instruction, so the full data dependency is kept (nothing is read or
// written before the predecessor grid has completed and flushed): only launch latency, CTA scheduling and the kernel
// prologue move under the predecessor's tail.  Measured slower (see pipeline.cu) and therefore opt-in
// (TK_B200_PDL_CHAIN=1); off (plain launches) for every other caller in any case.
bool& pdl_chain_hint();

// Number of kernels this library has enqueued (every launcher goes through
// launch_status); read by tk_b200_launch_count() for bench.py's gpu_launches.
void count_launch();

// Debug trace (tk_b200_debug_trace): when a device buffer is registered, the pre-process and post-process kernels
// append one 64-byte record per CTA (pre-process) or per frame (post-process) — SM id and %globaltimer stamps — which
// is how the co-residency timeline in profiles/ was taken (there is no nsys in this image).  One predicated store per
// CTA when enabled; nothing when the pointer is null.
struct TraceBuf { unsigned long long* buf; uint32_t cap; };   // buf[0]: records written; record i at buf[8 + 8 i]
TraceBuf trace_target();

// Post-launch check shared by every launcher (reference convention:
// cudaGetLastError after enqueue, src/gpu/cuda/tk_cuda_kernels.cu:297-300).
inline tk_error_code_t launch_status(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        tk_error_set_detail("%s launch failed: %s", what, cudaGetErrorString(e));
        return TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE;
    }
    count_launch();
    return TK_SUCCESS;
}

// CHECKED build (make -C trackiellm_b200/csrc checked -> libtk_b200_checked.so, selected with TK_B200_LIB): device-side
// assertions on every hand-computed shared-memory index / global offset, and — in the pre-process kernel — a byte-for-
// byte comparison of every ring slot a consumer is about to read (and has just read) with the source row in global
// memory, which turns any race of the producer / consumer protocol (a slot overwritten early, read before it landed,
// a phase mix-up) into a trap.  compute-sanitizer is closed on this GPU pool; this is what stands in for
// memcheck / racecheck (profiles/r2_sanitizer.md).  The shipped library is built without it.
#ifdef TK_B200_CHECKED
#define TK_DEV_ASSERT(cond)                                                                          \
    do {                                                                                             \
        if (!(cond)) {                                                                               \
            printf("TK_B200_CHECKED %s:%d block %u thread %u: %s\n", __FILE__, __LINE__, blockIdx.x, threadIdx.x, #cond); \
            __trap();                                                                                \
        }                                                                                            \
    } while (0)
#else
#define TK_DEV_ASSERT(cond) ((void)0)
#endif

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}
__device__ __forceinline__ uint32_t my_smid() { uint32_t s; asm volatile("mov.u32 %0, %%smid;" : "=r"(s)); return s; }
// record: {kernel id << 32 | a, smid << 32 | b, t[0..5]}
__device__ __forceinline__ void trace_put(const TraceBuf& t, uint32_t kid, uint32_t a, uint32_t b, const unsigned long long* ts, int nts) {
    const unsigned long long i = atomicAdd(t.buf, 1ull);
    if (i >= t.cap) return;
    unsigned long long* r = t.buf + 8 + 8 * i;
    r[0] = ((unsigned long long)kid << 32) | a;
    r[1] = ((unsigned long long)my_smid() << 32) | b;
    for (int j = 0; j < 6; ++j) r[2 + j] = j < nts ? ts[j] : 0ull;
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t addr = smem_u32(bar), ok;
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
// Same, with a sleep between polls so a waiting warp does not burn issue slots.
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
    uint32_t addr = smem_u32(bar), ok;
    for (;;) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (ok) break;
        __nanosleep(64);
    }
}
// Same, the barrier given by its shared-window address.
__device__ __forceinline__ void mbar_wait_backoff_addr(uint32_t addr, uint32_t parity) {
    uint32_t ok;
    for (;;) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
        if (ok) break;
        __nanosleep(64);
    }
}
// CTA-scope acquire load / release store on shared memory (flag hand-over between warps without a separate fence).
__device__ __forceinline__ uint32_t lds_acquire(const volatile uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(const_cast<const uint32_t*>(p))) : "memory");
    return v;
}
__device__ __forceinline__ void sts_release(volatile uint32_t* p, uint32_t v) {
    asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(smem_u32(const_cast<uint32_t*>(p))), "r"(v) : "memory");
}
// TMA 1-D bulk copy global -> shared, completion signalled on an mbarrier.
// dst, src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// 4-byte cp.async (LDGSTS): global -> shared with no register in between, any 4-byte alignment, also on page-locked host
// memory; completion by commit / wait groups of the issuing thread.
__device__ __forceinline__ void cp_async4(void* dst_smem, const void* src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- packed fp32 (sm_100: two IEEE round-to-nearest lanes per issued instruction) ----------------
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 f2s(float a) { return make_float2(a, a); }
// ptxas 12.9 contracts a packed multiply followed by a packed add into FFMA2 even for mul.rn.f32x2 /
// add.rn.f32x2 and under -fmad=false (seen in SASS; results then differ from the reference by an ulp).
// An add whose operand is a product is therefore written as sum2(): fma(x, 1.0, y) with the 1.0 read
// from the kernel parameters — one rounding of the exact sum, the same result as the add, nothing to contract.
__device__ __forceinline__ unsigned long long pk(float2 a) {
    unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y)); return r;
}
__device__ __forceinline__ float2 upk(unsigned long long r) {
    float2 a; asm("mov.b64 {%0, %1}, %2;" : "=f"(a.x), "=f"(a.y) : "l"(r)); return a;
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
    unsigned long long r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return upk(r);
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
    unsigned long long r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(pk(a)), "l"(pk(b))); return upk(r);
}
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
    unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(pk(a)), "l"(pk(b)), "l"(pk(c))); return upk(r);
}
__device__ __forceinline__ float2 sum2(float2 x, float2 y, float one) { return fma2(x, f2s(one), y); }
#endif

}  // namespace tkb
