// nms.cu — class-wise greedy IoU non-maximum suppression, one CTA per frame.
//
// Completes  the NMS step the reference leaves as a comment
//              (src/vision/tk_object_detector.c:357) with calculate_iou a stub
//              (:371), and replaces tk_cuda_nms_kernel / tk_cuda_tensor_nms
//              (src/gpu/extensions/cuda/tk_cuda_tensor_ops.cu:452-482, :858-889),
//              which only filter by score.
// Semantics:
//   IoU on truncated integer rects exactly as src/vision/src/object_analysis.rs:71-91;
//   order (confidence desc by IEEE total order, proposal index asc);
//   box k is dropped iff an already KEPT box j<k of the same class has IoU > thr
//   (strict; predicate of src/gpu/metal/tk_metal_kernels.metal:868-875 made sequential);
//   output = kept boxes in sorted order, truncated to max_detections.
//
// B200 design (latency-bound, filled by frame-level parallelism; shared memory only):
//   * candidates -> 64-bit keys (~order_key(conf) << 32 | index) + position payload,
//     bitonic sort in shared memory; boxes/classes staged in score order;
//   * a second bitonic sort on (class << 32 | score rank) makes every class one
//     contiguous segment with score order inside.  Suppression never crosses classes,
//     so each candidate builds ONE 64-bit mask of the same-class predecessors that
//     would suppress it (bitmask matrix, block-diagonal by class; neighbouring lanes
//     run equally long loops, no divergence), and the head of every segment resolves
//     its segment with a register-carried kept mask — all classes in parallel;
//   * kept bits are scattered back to score rank and compacted by ballot + prefix;
//   * a class segment longer than 64 (a crowd of one class, class-agnostic mode)
//     switches the frame to the blocked path: 64-column blocks of the suppression
//     matrix over the score-sorted list (diagonal block per thread, one-thread
//     resolve, CTA-wide sweep);
//   * more than 4096 candidates per frame (the shared-memory limit): the same nms_frame code with its work area in a
//     stream-ordered global scratch (nms_global_kernel) — slow, rare, and what keeps a low threshold or a crowded frame
//     from being silently truncated: the reference keeps every candidate (calloc(num_proposals),
//     src/vision/tk_object_detector.c:339-355).
#include "nms_frame.cuh"

#include <mutex>

namespace tkb {

// One CTA per frame while the batch fits one wave; larger batches loop (every exit of nms_frame is CTA-uniform).
__global__ void __launch_bounds__(kNmsThreads, 3) nms_kernel(const NmsK k) {
    // Nothing this grid reads is written by the kernel the pipeline launches next (the pre-process kernel of the
    // following step, which ends on griddepcontrol.wait): let it start on the SMs this grid does not fill.
    asm volatile("griddepcontrol.wait;" ::: "memory");            // (launched under decode_kernel's tail by the pipeline; else a no-op)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    extern __shared__ __align__(16) uint8_t dsm[];
    for (uint32_t frame = blockIdx.x; frame < k.num_frames; frame += gridDim.x) {
        nms_frame<kNmsThreads>(k, frame, dsm, k.Kp);
        __syncthreads();
    }
}

// Any number of candidates per frame: the work area of each CTA lives in global memory (L2-resident while it is hot).
__global__ void __launch_bounds__(kNmsThreads, 2) nms_global_kernel(const NmsK k, uint8_t* scratch, const uint32_t pitch) {
    uint8_t* work = scratch + (size_t)blockIdx.x * nms_smem_bytes(pitch);
    for (uint32_t frame = blockIdx.x; frame < k.num_frames; frame += gridDim.x) {
        nms_frame<kNmsThreads>(k, frame, work, pitch);
        __syncthreads();
    }
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_kernels_nms_classwise(const tk_b200_nms_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_counts || !p->d_num_detections || (!p->d_candidates && p->max_candidates) ||
        (!p->d_detections && p->max_detections))
        return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_frames == 0) return TK_SUCCESS;
    if (p->max_candidates > (1u << 30)) return TK_ERROR_INVALID_ARGUMENT;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    NmsK k{};
    k.cands = p->d_candidates; k.counts = p->d_counts; k.dets = p->d_detections; k.num_dets = p->d_num_detections;
    k.cap = p->max_candidates; k.max_det = p->max_detections;
    k.Kp = 512; while (k.Kp < k.cap) k.Kp <<= 1;    // >= 512: the register sort exchanges through 512-entry buffers
    k.thr = p->iou_threshold; k.agnostic = p->class_agnostic;
    const bool any_restore = p->restore_src_width || p->restore_src_height || p->restore_in_width || p->restore_in_height;
    if (any_restore) {
        if (!p->restore_src_width || !p->restore_src_height || !p->restore_in_width || !p->restore_in_height)
            return TK_ERROR_INVALID_DIMENSIONS;
        k.restore = 1;
        k.sx = (float)p->restore_src_width / (float)p->restore_in_width;
        k.sy = (float)p->restore_src_height / (float)p->restore_in_height;
    }
    k.num_frames = p->num_frames;
    k.trace = trace_target();
    if (p->max_candidates > kNmsMaxK) {
        // above the shared-memory layout: global work areas from the stream-ordered allocator, one per CTA
        uint32_t pitch = 2 * kNmsMaxK;
        while (pitch < p->max_candidates) pitch <<= 1;
        const uint32_t grid = p->num_frames < 2u * (uint32_t)dev->sm_count ? p->num_frames : 2u * (uint32_t)dev->sm_count;
        uint8_t* scratch = nullptr;
        if (cudaMallocAsync(reinterpret_cast<void**>(&scratch), (size_t)grid * nms_smem_bytes(pitch), stream) != cudaSuccess) {
            tk_error_set_detail("nms: %zu B of scratch for %u candidates per frame: %s", (size_t)grid * nms_smem_bytes(pitch),
                                p->max_candidates, cudaGetErrorString(cudaGetLastError()));
            return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
        }
        nms_global_kernel<<<grid, kNmsThreads, 0, stream>>>(k, scratch, pitch);
        const tk_error_code_t rc = launch_status("nms_global_kernel");
        cudaFreeAsync(scratch, stream);
        return rc;
    }
    const size_t smem = nms_smem_bytes(k.Kp);
    if (smem > 48 * 1024)
        TK_CUDA_TRY(cudaFuncSetAttribute(nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), TK_ERROR_GPU_CUDA_ERROR);
    const uint32_t sms = (uint32_t)dev->sm_count;
    if (p->num_frames <= sms || k.Kp <= kNmsSmall) {
        // one wave (or the small layout anyway): a single launch sized for the cap
        int per_sm = 1;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nms_kernel, kNmsThreads, smem) != cudaSuccess || per_sm < 1) {
            cudaGetLastError(); per_sm = 1;
        }
        const uint32_t cap = sms * (uint32_t)per_sm;
        if (pdl_chain_hint()) {
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(p->num_frames < cap ? p->num_frames : cap); cfg.blockDim = dim3(kNmsThreads);
            cfg.dynamicSmemBytes = smem; cfg.stream = stream;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            TK_CUDA_TRY(cudaLaunchKernelEx(&cfg, nms_kernel, k), TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE);
        } else {
            nms_kernel<<<p->num_frames < cap ? p->num_frames : cap, kNmsThreads, smem, stream>>>(k);
        }
        return launch_status("nms_kernel");
    }
    // More frames than SMs: shared memory sized for the cap (200 KB at 4096 candidates) means one CTA per SM and
    // batch / 148 waves.  Frames of <= 512 candidates — the normal case — only use the register path, which needs
    // 25 KB: they go first, several CTAs per SM; a second one-wave launch takes the crowded frames and skips the rest.
    NmsK small = k;
    small.Kp = kNmsSmall; small.select = 1;
    const size_t smem_small = nms_smem_bytes(kNmsSmall);
    int per_sm = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, nms_kernel, kNmsThreads, smem_small) != cudaSuccess || per_sm < 1) {
        cudaGetLastError(); per_sm = 1;
    }
    const uint32_t cap_small = sms * (uint32_t)per_sm;
    nms_kernel<<<p->num_frames < cap_small ? p->num_frames : cap_small, kNmsThreads, smem_small, stream>>>(small);
    tk_error_code_t rc = launch_status("nms_kernel");
    if (rc != TK_SUCCESS) return rc;
    k.select = 2;
    nms_kernel<<<sms, kNmsThreads, smem, stream>>>(k);
    return launch_status("nms_kernel");
}

// ------------------------------------------------------------------------------------------------------------------ //
// tk_b200_nms_classwise_static — the same NMS with NO allocation of any kind (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The per-CTA work areas of nms_global_kernel live in a file-scope device array
// instead of stream-ordered scratch.  Every GPU that loads the library holds the array (kNmsStaticBytes, 6.1 MiB) until
// the process exits; every call on a device shares it, so calls that overlap on the GPU (two streams, two host
// threads) race — serialise them.  The array keeps the last call's contents: nms_frame initialises every word it
// reads (keys, boxes, masks, the kept bitmap) per frame, nothing has to start at zero.
// ------------------------------------------------------------------------------------------------------------------ //
namespace tkb {
static constexpr uint32_t kNmsStaticMaxK = 16384;                              // most candidates per frame this entry point takes
static constexpr uint32_t kNmsStaticAreas = 8;                                 // work areas at that size (more when frames are smaller)
static constexpr size_t kNmsStaticBytes = (size_t)kNmsStaticAreas * ((size_t)kNmsStaticMaxK * 49 + kNmsStaticMaxK / 8);
static_assert(kNmsStaticBytes == kNmsStaticAreas * nms_smem_bytes(kNmsStaticMaxK), "work-area layout of nms_frame.cuh");
static __device__ __align__(256) uint8_t g_nms_static_scratch[kNmsStaticBytes];

// Device address of the array on the CURRENT device, looked up once per device (in host code the array's name is a
// host-side stand-in, not something a kernel can be handed).
static tk_error_code_t nms_static_scratch(uint8_t** out) {
    static std::mutex mu;
    static uint8_t* addr[64] = {};
    int dev = -1;
    if (!device_info(&dev)) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    std::lock_guard<std::mutex> lk(mu);
    if (!addr[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_nms_static_scratch);
        if (e != cudaSuccess || !ptr) {
            cudaGetLastError();
            tk_error_set_detail("nms (static scratch): cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return TK_ERROR_GPU_CUDA_ERROR;
        }
        addr[dev] = static_cast<uint8_t*>(ptr);
    }
    *out = addr[dev];
    return TK_SUCCESS;
}
}  // namespace tkb

extern "C" uint32_t tk_b200_nms_classwise_static_max_candidates(void) { return tkb::kNmsStaticMaxK; }

extern "C" tk_error_code_t tk_b200_nms_classwise_static(const tk_b200_nms_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_counts || !p->d_num_detections || (!p->d_candidates && p->max_candidates) ||
        (!p->d_detections && p->max_detections))
        return TK_ERROR_INVALID_ARGUMENT;
    if (p->max_candidates > kNmsStaticMaxK) {
        tk_error_set_detail("nms (static scratch): max_candidates %u is above the %u this entry point holds scratch for; "
                            "use tk_kernels_nms_classwise", p->max_candidates, kNmsStaticMaxK);
        return TK_ERROR_BUFFER_TOO_SMALL;
    }
    if (p->num_frames == 0) return TK_SUCCESS;
    NmsK k{};
    k.cands = p->d_candidates; k.counts = p->d_counts; k.dets = p->d_detections; k.num_dets = p->d_num_detections;
    k.cap = p->max_candidates; k.max_det = p->max_detections;
    k.Kp = 512; while (k.Kp < k.cap) k.Kp <<= 1;    // >= 512: the register sort exchanges through 512-entry buffers
    k.thr = p->iou_threshold; k.agnostic = p->class_agnostic;
    const bool any_restore = p->restore_src_width || p->restore_src_height || p->restore_in_width || p->restore_in_height;
    if (any_restore) {
        if (!p->restore_src_width || !p->restore_src_height || !p->restore_in_width || !p->restore_in_height)
            return TK_ERROR_INVALID_DIMENSIONS;
        k.restore = 1;
        k.sx = (float)p->restore_src_width / (float)p->restore_in_width;
        k.sy = (float)p->restore_src_height / (float)p->restore_in_height;
    }
    k.num_frames = p->num_frames;
    uint8_t* scratch = nullptr;
    const tk_error_code_t rc = nms_static_scratch(&scratch);
    if (rc != TK_SUCCESS) return rc;
    const uint32_t pitch = k.Kp;                                               // one work area = nms_smem_bytes(pitch)
    const size_t areas = kNmsStaticBytes / nms_smem_bytes(pitch);              // 8 at 16384 candidates ... 256 at 512
    const uint32_t grid = p->num_frames < areas ? p->num_frames : (uint32_t)areas;
    nms_global_kernel<<<grid, kNmsThreads, 0, stream>>>(k, scratch, pitch);
    return launch_status("nms_global_kernel (static scratch)");
}
