// post.cu — co-resident post-process: YOLO-head decode + class-wise NMS of a frame in ONE persistent kernel.
//
// Completes  postprocess_detections (src/vision/tk_object_detector.c:303-368): the candidate loop :343-355
//            (decode_row.cuh) and the NMS step the reference leaves as a comment (:357; nms_frame.cuh).
//
// Why a second form of decode_kernel + nms_kernel: in a pipeline step those two run behind the pre-process kernel
// although they share no data with it (logits vs frames), and both are latency-bound — decode_kernel is sized to
// fill the machine (64 clusters x 8 CTAs x 512 threads) and nms_kernel needs 200 KB of shared memory per CTA, so
// neither can be resident beside the pre-process kernel's two CTAs per SM (2 x 96 KB of shared memory, 2 x 352
// threads x 72-80 registers).  This kernel is cut to fit in what those leave free on every SM —
//     256 threads, <= 32 registers, 25 KB of dynamic shared memory
// — and is launched on a second stream, so the whole post-process of a batch runs UNDER the pre-process kernel
// instead of after it.  One persistent CTA takes whole frames (frame = blockIdx.x, += gridDim.x):
//   pass 1   objectness word of every row (ld.global.nc.L2::64B, 8 loads in flight per thread) -> warp-ballot bitmap
//            in shared memory -> exclusive prefix (one warp): survivor s knows its slot, ascending proposal order
//   pass 2   surviving rows are pulled into shared memory with 4-byte cp.async copies (no registers, any alignment,
//            works on rows in page-locked host memory): every warp keeps two groups of R rows in flight (R = 3 at
//            85-word rows: 48 rows per CTA), decodes a group while the next one lands, and writes the 28-byte records
//            to the frame's candidate list in global memory (L2-resident; tk_b200_pipeline keeps that buffer anyway)
//   NMS      nms_frame<256> on the list: in shared memory when the frame has <= 512 candidates (the decode-phase
//            buffers are dead by then and are overlaid), else through a per-CTA scratch in global memory — same code,
//            ordinary loads / stores / atomics between CTA barriers — so crowded frames are slow, not wrong.
// Results are bit-identical to decode_kernel + nms_kernel: same decode_row, same nms_frame, same candidate list.
#include "decode_row.cuh"
#include "nms_frame.cuh"

namespace tkb {

static constexpr int kPostThreads = 256;
static constexpr int kPostWarps = kPostThreads / 32;
static constexpr int kPostUnroll = 8;             // objectness loads in flight per thread
static constexpr uint32_t kPostList = 512;        // surviving-row numbers kept in shared memory (later ones: read back from the list)
static constexpr uint32_t kPostMaxStage = 4;      // rows per warp per staging buffer, at most

struct PostK {
    DecK d;                  // rows_per_cta = P rounded up to 32 (one CTA owns the frame), nwords = rows_per_cta / 32
    NmsK n;
    uint8_t* scratch;        // [gridDim.x][nms_smem_bytes(scratch_pitch)]: NMS work area of frames above kNmsSmall candidates
    uint32_t scratch_pitch;  // entries (power of two >= cap), 0 when cap <= kNmsSmall
    uint32_t stage_rows;     // R; 0 = rows are decoded straight from global memory
    TraceBuf trace;          // debug: per-frame record (tk_b200_debug_trace)
};

template <int ACT, bool PACKED>
__global__ void __launch_bounds__(kPostThreads, 8) post_kernel(const PostK k) {
    extern __shared__ __align__(16) uint8_t dsm[];
    __shared__ uint32_t s_total;
    const DecK& d = k.d;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // decode-phase layout of the dynamic shared memory (the NMS phase overlays it)
    uint32_t* flags = reinterpret_cast<uint32_t*>(dsm);            // [nwords] ballot bitmap
    uint32_t* wpre = flags + d.nwords;                             // [nwords] exclusive prefix of popc
    uint32_t* rowlist = wpre + d.nwords;                           // [kPostList] row number of survivor s
    float* stage = reinterpret_cast<float*>(rowlist + kPostList);  // [warps][2][R][L] staged rows (R == 0: [warps][C] softmax scratch)
    const uint32_t R = k.stage_rows;

    for (uint32_t frame = blockIdx.x; frame < k.n.num_frames; frame += gridDim.x) {
        const float* fl = d.logits + (size_t)frame * d.P * d.L;
        unsigned long long tt[5];
        const bool tr = k.trace.buf && tid == 0;
        if (tr) tt[0] = gtimer();
        // ---- pass 1: objectness threshold, ballot bitmap -------------------------------------------------
        for (uint32_t base = 0; base < d.rows_per_cta; base += kPostUnroll * kPostThreads) {
            float obj[kPostUnroll];
#pragma unroll
            for (int u = 0; u < kPostUnroll; ++u) {
                const uint32_t r = base + u * kPostThreads + tid;
                if constexpr (PACKED) obj[u] = r < d.P ? __ldg(d.obj + (size_t)frame * d.P + r) : __int_as_float(0x7fc00000);
                else obj[u] = r < d.P ? ld_obj(fl + (size_t)r * d.L + 4) : __int_as_float(0x7fc00000);
            }
#pragma unroll
            for (int u = 0; u < kPostUnroll; ++u) {
                const uint32_t r = base + u * kPostThreads + tid;
                const bool pass = act_fn<ACT>(obj[u]) > d.thr;     // NaN never passes
                const uint32_t b = __ballot_sync(0xffffffffu, pass);
                if (lane == 0 && r < d.rows_per_cta) flags[r >> 5] = b;
            }
        }
        __syncthreads();
        if (tr) tt[1] = gtimer();
        // ---- exclusive prefix over the bitmap words (one warp) --------------------------------------------
        if (warp == 0) {
            uint32_t run = 0;
            for (uint32_t w0 = 0; w0 < d.nwords; w0 += 32) {
                const uint32_t w = w0 + lane;
                const uint32_t c = w < d.nwords ? __popc(flags[w]) : 0u;
                uint32_t inc = c;
#pragma unroll
                for (int s = 1; s < 32; s <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, s);
                    if ((int)lane >= s) inc += t;
                }
                if (w < d.nwords) wpre[w] = run + inc - c;
                run += __shfl_sync(0xffffffffu, inc, 31);
            }
            if (lane == 0) { s_total = run; d.counts[frame] = run; }
        }
        __syncthreads();
        const uint32_t n_out = min(s_total, d.cap);
        tk_b200_candidate_t* out = d.cands + (size_t)frame * d.cap;
        // survivor s -> row number: thread per bit; the first kPostList in shared memory, later ones into the index
        // word of their record (where the index goes anyway)
        for (uint32_t i = tid; i < d.nwords * 32; i += kPostThreads) {
            const uint32_t word = flags[i >> 5], b = i & 31;
            if ((word >> b) & 1u) {
                const uint32_t lpos = wpre[i >> 5] + __popc(word & ((1u << b) - 1u));
                if (lpos < kPostList) rowlist[lpos] = i;
                else if (lpos < n_out) out[lpos].index = (int32_t)i;
            }
        }
        __syncthreads();
        if (tr) tt[2] = gtimer();
        // ---- pass 2: this warp's survivors are warp, warp + 8, ... -------------------------------------------
        auto row_of = [&](uint32_t sidx) -> uint32_t { return sidx < kPostList ? rowlist[sidx] : (uint32_t)out[sidx].index; };
        const uint32_t mine = n_out > warp ? (n_out - warp + kPostWarps - 1) / kPostWarps : 0u;
        if (R) {
            float* wst = stage + (size_t)warp * 2 * R * d.L;
            auto issue = [&](uint32_t g) {
                for (uint32_t i = 0; i < R; ++i) {
                    const uint32_t j = g * R + i;
                    if (j < mine) {
                        const float* src = fl + (size_t)row_of(warp + kPostWarps * j) * d.L;
                        float* dst = wst + (size_t)((g & 1u) * R + i) * d.L;
                        for (uint32_t w = lane; w < d.L; w += 32) cp_async4(dst + w, src + w);
                    }
                }
                cp_async_commit();
            };
            const uint32_t G = (mine + R - 1) / R;
            if (G) issue(0);
            for (uint32_t g = 0; g < G; ++g) {
                if (g + 1 < G) { issue(g + 1); cp_async_wait<1>(); } else cp_async_wait<0>();
                __syncwarp();
                for (uint32_t i = 0; i < R; ++i) {
                    const uint32_t j = g * R + i;
                    if (j >= mine) break;
                    const uint32_t sidx = warp + kPostWarps * j, r = row_of(sidx);
                    TK_DEV_ASSERT(sidx < n_out && r < d.P);
                    float* row = wst + (size_t)((g & 1u) * R + i) * d.L;
                    TK_DEV_ASSERT(reinterpret_cast<uint8_t*>(row + d.L) <= dsm + nms_smem_bytes(kNmsSmall));
                    RowRegs q;
                    load_row<false>(d, row, lane, q);
                    // softmax: the exponentials overwrite the staged class logits in place (each lane reads word j before it writes it)
                    const tk_b200_candidate_t c = decode_row<ACT, false>(d, row, r, lane, row + 5, q);
                    if (lane == 0) out[sidx] = c;
                }
                __syncwarp();                                      // the buffer is reused two groups later
            }
        } else {
            float* my_e = stage + (size_t)warp * d.C;
            for (uint32_t j = 0; j < mine; ++j) {
                const uint32_t sidx = warp + kPostWarps * j, r = row_of(sidx);
                RowRegs q;
                load_row<true>(d, fl + (size_t)r * d.L, lane, q);
                const tk_b200_candidate_t c = decode_row<ACT, true>(d, fl + (size_t)r * d.L, r, lane, my_e, q);
                if (lane == 0) out[sidx] = c;
            }
        }
        __syncthreads();               // the frame's candidate list and count are in global memory, visible to this CTA
        if (tr) tt[3] = gtimer();
        // ---- NMS ----------------------------------------------------------------------------------------
        if (n_out <= kNmsSmall) nms_frame<kPostThreads>(k.n, frame, dsm, kNmsSmall);
        else nms_frame<kPostThreads>(k.n, frame, k.scratch + (size_t)blockIdx.x * nms_smem_bytes(k.scratch_pitch), k.scratch_pitch);
        __syncthreads();
        if (tr) { tt[4] = gtimer(); trace_put(k.trace, 2u, blockIdx.x, frame, tt, 5); }
    }
}

template <int ACT>
static tk_error_code_t launch_post(const PostK& k, uint32_t grid, size_t smem, cudaStream_t stream) {
    auto kern = k.d.obj ? post_kernel<ACT, true> : post_kernel<ACT, false>;
    // The point of this kernel is to sit beside another kernel's CTAs: ask for the largest shared-memory carve-out so
    // the SM never has to drain to change its L1 / shared split between the two.
    static bool once = [&] {
        cudaFuncSetAttribute(post_kernel<ACT, true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaFuncSetAttribute(post_kernel<ACT, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        cudaGetLastError();
        return true;
    }();
    (void)once;
    kern<<<grid, kPostThreads, smem, stream>>>(k);
    return launch_status("post_kernel");
}

static uint32_t post_scratch_pitch(uint32_t max_candidates) {
    uint32_t pitch = 1024;
    while (pitch < max_candidates) pitch <<= 1;
    return pitch;
}
// layout 0 (cluster per frame): one work area per frame; layout 1 (persistent CTAs): one per CTA, at most one per SM
size_t post_fused_workspace_bytes(uint32_t max_candidates, uint32_t num_frames, int layout, int sm_count) {
    if (max_candidates <= kNmsSmall) return 0;
    const uint32_t areas = layout == 0 ? num_frames : (num_frames < (uint32_t)sm_count ? num_frames : (uint32_t)sm_count);
    return (size_t)areas * nms_smem_bytes(post_scratch_pitch(max_candidates));
}

}  // namespace tkb

using namespace tkb;

extern "C" size_t tk_b200_postprocess_fused_workspace_bytes(uint32_t max_candidates, uint32_t num_frames, int layout) {
    const DeviceInfo* dev = device_info();
    return dev ? post_fused_workspace_bytes(max_candidates, num_frames, layout, dev->sm_count) : 0;
}

tk_error_code_t tkb::postprocess_fused_packed(const tk_b200_post_params_t* p, const float* d_objectness, cudaStream_t stream) {
    if (!p || !p->d_logits || !p->d_counts || !p->d_num_detections || (!p->d_candidates && p->max_candidates) ||
        (!p->d_detections && p->max_detections))
        return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_frames == 0) return TK_SUCCESS;
    if (p->num_proposals == 0) return TK_ERROR_INVALID_DIMENSIONS;      // dims check of tk_object_detector.c:325-336
    if ((unsigned)p->activation > TK_B200_ACT_SOFTMAX) return TK_ERROR_INVALID_ARGUMENT;
    if (p->max_candidates > (1u << 30)) return TK_ERROR_INVALID_ARGUMENT;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    PostK k{};
    k.d.logits = p->d_logits; k.d.obj = d_objectness; k.d.cands = p->d_candidates; k.d.counts = p->d_counts;
    k.d.P = p->num_proposals; k.d.C = p->num_classes; k.d.L = 5 + p->num_classes;
    k.d.rows_per_cta = (k.d.P + 31) & ~31u;
    k.d.nwords = k.d.rows_per_cta / 32;
    k.d.cap = p->max_candidates;
    k.d.thr = p->confidence_threshold;
    k.n.cands = p->d_candidates; k.n.counts = p->d_counts; k.n.dets = p->d_detections; k.n.num_dets = p->d_num_detections;
    k.n.cap = p->max_candidates; k.n.max_det = p->max_detections;
    k.n.thr = p->iou_threshold; k.n.agnostic = p->class_agnostic;
    const bool any_restore = p->restore_src_width || p->restore_src_height || p->restore_in_width || p->restore_in_height;
    if (any_restore) {
        if (!p->restore_src_width || !p->restore_src_height || !p->restore_in_width || !p->restore_in_height)
            return TK_ERROR_INVALID_DIMENSIONS;
        k.n.restore = 1;
        k.n.sx = (float)p->restore_src_width / (float)p->restore_in_width;
        k.n.sy = (float)p->restore_src_height / (float)p->restore_in_height;
    }
    k.n.num_frames = p->num_frames;
    if ((unsigned)p->layout > 1u) return TK_ERROR_INVALID_ARGUMENT;
    const uint32_t grid = p->num_frames < (uint32_t)dev->sm_count ? p->num_frames : (uint32_t)dev->sm_count;
    if (p->max_candidates > kNmsSmall) {
        const size_t need = post_fused_workspace_bytes(p->max_candidates, p->num_frames, p->layout, dev->sm_count);
        if (!p->d_workspace || p->workspace_bytes < need || (reinterpret_cast<uintptr_t>(p->d_workspace) & 15u)) {
            tk_error_set_detail("post: workspace of %zu bytes (16-byte aligned) needed for max_candidates %u", need, p->max_candidates);
            return TK_ERROR_BUFFER_TOO_SMALL;
        }
        k.scratch = static_cast<uint8_t*>(p->d_workspace);
        k.scratch_pitch = post_scratch_pitch(p->max_candidates);
    }
    if (p->layout == 0) {
        // a cluster of 8 CTAs per frame: decode_kernel with the NMS tail (decode.cu)
        tk_b200_decode_params_t dp{};
        dp.d_logits = p->d_logits; dp.num_frames = p->num_frames; dp.num_proposals = p->num_proposals; dp.num_classes = p->num_classes;
        dp.confidence_threshold = p->confidence_threshold; dp.activation = p->activation;
        dp.d_candidates = p->d_candidates; dp.d_counts = p->d_counts; dp.max_candidates = p->max_candidates;
        return decode_nms_cluster(&dp, d_objectness, &k.n, k.scratch, k.scratch_pitch, stream);
    }
    // Dynamic shared memory: exactly the small NMS layout (25 KB) — what fits beside two pre-process CTAs.  The decode
    // phase takes its bitmap, prefix and row list out of it and stages rows in the rest.
    const size_t smem = nms_smem_bytes(kNmsSmall);
    const size_t fixed = (size_t)k.d.nwords * 8 + kPostList * 4;
    if (fixed > smem) {
        tk_error_set_detail("post: %u proposals need %zu B of shared memory for the bitmap", k.d.P, fixed);
        return TK_ERROR_NOT_IMPLEMENTED;                               // callers fall back to decode_kernel + nms_kernel
    }
    const size_t avail = smem - fixed;
    uint32_t R = (uint32_t)(avail / ((size_t)kPostWarps * 2 * k.d.L * 4));
    if (R > kPostMaxStage) R = kPostMaxStage;
    if (R == 0 && p->activation == TK_B200_ACT_SOFTMAX && avail < (size_t)kPostWarps * k.d.C * 4) {
        tk_error_set_detail("post: %u classes do not fit the softmax scratch", k.d.C);
        return TK_ERROR_NOT_IMPLEMENTED;
    }
    k.stage_rows = R;
    k.trace = trace_target();
    switch (p->activation) {
        case TK_B200_ACT_RAW:     return launch_post<TK_B200_ACT_RAW>(k, grid, smem, stream);
        case TK_B200_ACT_SIGMOID: return launch_post<TK_B200_ACT_SIGMOID>(k, grid, smem, stream);
        default:                  return launch_post<TK_B200_ACT_SOFTMAX>(k, grid, smem, stream);
    }
}

// Decode + class-wise NMS of ONE frame with NO candidate cap: the reference keeps every row that passes the threshold
// (calloc(num_proposals), src/vision/tk_object_detector.c:339-355).  The composed entry points (detect,
// tk_b200_postprocess_detections) run the capped kernels first and come here only for a frame whose count exceeded the
// cap: decode with room for all P rows, NMS through nms_global_kernel.  Stream-ordered scratch; asynchronous.
tk_error_code_t tkb::postprocess_frame_uncapped(const float* d_logits, uint32_t P, uint32_t C, tk_b200_activation_e act,
                                               float conf_thr, float iou_thr, uint32_t max_det, int class_agnostic,
                                               const uint32_t restore[4], tk_b200_candidate_t* d_dets, uint32_t* d_ndets,
                                               cudaStream_t st) {
    uint8_t* buf = nullptr;
    const size_t cand_bytes = ((size_t)P * sizeof(tk_b200_candidate_t) + 255) & ~(size_t)255;
    if (cudaMallocAsync(reinterpret_cast<void**>(&buf), cand_bytes + 256, st) != cudaSuccess) {
        tk_error_set_detail("uncapped post-process: %zu B: %s", cand_bytes + 256, cudaGetErrorString(cudaGetLastError()));
        return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
    }
    tk_b200_candidate_t* d_c = reinterpret_cast<tk_b200_candidate_t*>(buf);
    uint32_t* d_n = reinterpret_cast<uint32_t*>(buf + cand_bytes);
    tk_b200_decode_params_t d{};
    d.d_logits = d_logits; d.num_frames = 1; d.num_proposals = P; d.num_classes = C; d.confidence_threshold = conf_thr;
    d.activation = act; d.d_candidates = d_c; d.d_counts = d_n; d.max_candidates = P;
    tk_error_code_t rc = tk_kernels_decode_detections(&d, st);
    if (rc == TK_SUCCESS) {
        tk_b200_nms_params_t m{};
        m.d_candidates = d_c; m.d_counts = d_n; m.num_frames = 1; m.max_candidates = P; m.iou_threshold = iou_thr;
        m.max_detections = max_det; m.class_agnostic = class_agnostic; m.d_detections = d_dets; m.d_num_detections = d_ndets;
        if (restore) {
            m.restore_src_width = restore[0]; m.restore_src_height = restore[1];
            m.restore_in_width = restore[2]; m.restore_in_height = restore[3];
        }
        rc = tk_kernels_nms_classwise(&m, st);
    }
    cudaFreeAsync(buf, st);
    return rc;
}

extern "C" tk_error_code_t tk_kernels_postprocess_fused(const tk_b200_post_params_t* p, cudaStream_t stream) {
    return tkb::postprocess_fused_packed(p, nullptr, stream);
}
