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
//     the first-max class by shuffle reduction, and stages the 28-byte record in
//     shared memory in row order; the cluster barrier is split (arrive before
//     pass 2, wait after it) so the exchange hides behind the row decode, and the
//     staged records leave as one coalesced word copy.
//   * Sigmoid / softmax use tk_dm_expf (include/tk_b200_detmath.h), the same
//     source the oracle compiles, so thresholds and scores agree bit for bit.
#include "decode_row.cuh"
#include "nms_frame.cuh"

#include <stdlib.h>
#include <cooperative_groups.h>
namespace cg = cooperative_groups;

namespace tkb {

// NMS tail of the fused form (tk_kernels_postprocess_fused, layout 0): once the eight CTAs of a frame's cluster have
// written the candidate list, the first of them runs the class-wise NMS of that frame at once — in the shared memory
// the decode phase no longer needs when the frame has <= 512 candidates, else through the frame's global scratch —
// while the clusters of other frames are still decoding.  One launch instead of two, no gap between them, and the
// NMS of the early frames hides under the decode tail of the late ones.
struct NmsTail {
    NmsK n;
    uint8_t* scratch;        // [frames][nms_smem_bytes(scratch_pitch)], frames above kNmsSmall candidates
    uint32_t scratch_pitch;
};

// 4 CTAs per SM (<= 32 registers): with 3, only 48 of the 64 clusters of the bench batch are
// co-resident (a cluster of 8 must sit in one GPC) and the launch needs two waves.
// The form with the NMS tail: clusters of 4 and 3 CTAs per SM (<= 42 registers, which nms_frame's 40 fit without the
// spills it has at 32): 64 frames x 4 = 256 CTAs are still one wave.
static constexpr int kDecClusterNms = 4;

template <int ACT, bool PACKED, bool NMS>
__global__ void __launch_bounds__(kDecThreads, NMS ? 3 : 4) decode_kernel(const DecK k, const NmsTail tail) {
    constexpr int CL = NMS ? kDecClusterNms : kDecCluster;
    // Possibly launched under the tail of the pipeline's pre-process kernel (pdl_chain_hint): nothing is read or written
    // before that grid has completed.  A no-op for a plain launch.  nms_kernel, which starts the same way, is let in
    // when this CTA has written its records (end of the kernel): its launch then overlaps this grid's drain.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    extern __shared__ __align__(16) uint8_t dsm[];
    uint32_t* flags = reinterpret_cast<uint32_t*>(dsm);            // [nwords] ballot bitmap
    uint32_t* wpre = flags + k.nwords;                             // [nwords] exclusive prefix of popc
    float* ebuf = reinterpret_cast<float*>(wpre + k.nwords);       // softmax only: [warps][C]
    __shared__ uint32_t cta_tot[CL];                               // filled by every CTA of the cluster
    __shared__ uint32_t stage[kDecStage * 7];                      // this CTA's first kDecStage records, in row order
    __shared__ unsigned long long tstamp[6];                       // debug trace: start, pass 1, prefix, expand, pass 2, end
#ifdef TK_B200_TRACE_POST
    const bool tr = k.trace.buf != nullptr && threadIdx.x == 0;
#else
    constexpr bool tr = false;                                     // phase stamps: `make trace` builds only
#endif
    if (tr) tstamp[0] = gtimer();

    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t rank = cluster.block_rank();
    const uint32_t frame = blockIdx.x / CL;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* fl = k.logits + (size_t)frame * k.P * k.L;
    const uint32_t row0 = rank * k.rows_per_cta;

    // ---- pass 1: objectness threshold, ballot bitmap ------------------------
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");   // "I am resident" (waited on before DSMEM)
    if (k.bitmap) {
        // pass 1 was done by the scanner warps of the pre-process kernel: this CTA's slice of the frame's ballot words
        const uint32_t wbase = row0 >> 5;                          // row0 is a multiple of 32
        for (uint32_t i = tid; i < k.nwords; i += kDecThreads)
            flags[i] = wbase + i < k.nwf ? k.bitmap[(size_t)frame * k.nwf + wbase + i] : 0u;
    } else
    for (uint32_t base = 0; base < k.rows_per_cta; base += kDecUnroll * kDecThreads) {
        float obj[kDecUnroll];
#pragma unroll
        for (int u = 0; u < kDecUnroll; ++u) {
            const uint32_t lr = base + u * kDecThreads + tid, r = row0 + lr;
            if constexpr (PACKED)
                obj[u] = (lr < k.rows_per_cta && r < k.P) ? __ldg(k.obj + (size_t)frame * k.P + r) : __int_as_float(0x7fc00000);
            else
                obj[u] = (lr < k.rows_per_cta && r < k.P) ? ld_obj(fl + (size_t)r * k.L + 4) : __int_as_float(0x7fc00000);
        }
#pragma unroll
        for (int u = 0; u < kDecUnroll; ++u) {
            const uint32_t lr = base + u * kDecThreads + tid;
            const bool pass = act_fn<ACT>(obj[u]) > k.thr;         // NaN never passes
            if constexpr (PACKED) {
                // the rows are in host memory: start pulling a surviving row over PCIe now, pass 2 reads it a few
                // microseconds later.  (For rows in HBM the prefetch buys nothing now that pass 2 issues a row's loads
                // together: 0.5 - 1 us per 64 frames FASTER without it.)
                if (pass) {
                    const char* rp = reinterpret_cast<const char*>(fl + (size_t)(row0 + lr) * k.L);
                    const uint32_t rb = k.L * 4;
                    for (uint32_t o = 0; o < rb; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(rp + o));
                }
            }
            const uint32_t b = __ballot_sync(0xffffffffu, pass);
            if (lane == 0 && lr < k.rows_per_cta) flags[lr >> 5] = b;
        }
    }
    __syncthreads();
    if (tr) tstamp[1] = gtimer();
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");             // every peer CTA has started

    // ---- exclusive prefix over the bitmap words; totals to every peer through DSMEM ----
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
        if (lane < CL)
            *cluster.map_shared_rank(&cta_tot[rank], lane) = run;  // DSMEM store into every peer
    }
    __syncthreads();
    if (tr) tstamp[2] = gtimer();
    // The frame-level offset is only needed to write the records out: arrive now, wait after
    // the rows are decoded, so the cluster exchange hides behind pass 2.
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");

    // ---- pass 2: one warp per surviving row, records staged in shared memory in row order ----
    // Survivor s of this CTA goes to warp s mod 16 (balanced however the survivors cluster).  Its row number is
    // expanded from the bitmap once, thread per bit, into the first word of its record slot (where the index goes
    // anyway): pass 2 is issue-bound with 64 warps per SM in it at once, and a per-row binary search + select-nth-bit
    // by all 32 lanes cost more instructions than decoding the row.
    float* my_e = ebuf + (size_t)warp * k.C;
    const uint32_t local_n = __popc(flags[k.nwords - 1]) + wpre[k.nwords - 1];
    const uint32_t lim = min(local_n, kDecStage);
    for (uint32_t i = tid; i < k.nwords * 32; i += kDecThreads) {
        const uint32_t word = flags[i >> 5], b = i & 31;
        if ((word >> b) & 1u) {
            const uint32_t lpos = wpre[i >> 5] + __popc(word & ((1u << b) - 1u));
            if (lpos < kDecStage) stage[lpos * 7] = row0 + i;
        }
    }
    __syncthreads();
    if (tr) tstamp[3] = gtimer();
    auto row_of = [&](uint32_t sidx) -> uint32_t { return stage[sidx * 7]; };
    auto put = [&](uint32_t sidx, const tk_b200_candidate_t& c) {
        if (lane == 0) {
            uint32_t* sp = stage + sidx * 7;
            sp[0] = (uint32_t)c.index; sp[1] = (uint32_t)c.bbox.x; sp[2] = (uint32_t)c.bbox.y; sp[3] = (uint32_t)c.bbox.w;
            sp[4] = (uint32_t)c.bbox.h; sp[5] = __float_as_uint(c.confidence); sp[6] = c.class_id;
        }
    };
    // (Fetching all of a CTA's ~60 surviving rows into shared memory at once with cp.async — one DRAM round trip instead
    // of ~4 dependent ones per warp — was built and measured: 36.2 us against 35.0 us for 64 frames.  The pass is not a
    // latency chain: with 64 warps per SM in it, other warps' rows are in flight while one waits.)
    for (uint32_t sidx = warp; sidx < lim; sidx += kDecWarps) {
        const uint32_t r = row_of(sidx);
        RowRegs q;
        load_row(k, fl + (size_t)r * k.L, lane, q);
        put(sidx, decode_row<ACT>(k, fl + (size_t)r * k.L, r, lane, my_e, q));
    }
    __syncthreads();
    if (tr) tstamp[4] = gtimer();
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    uint32_t frame_pre = 0, frame_tot = 0;
#pragma unroll
    for (int r = 0; r < CL; ++r) {
        uint32_t t = cta_tot[r];
        frame_pre += (uint32_t)r < rank ? t : 0u;
        frame_tot += t;
    }
    if (rank == 0 && tid == 0) k.counts[frame] = frame_tot;
    const uint32_t local_tot = cta_tot[rank];

    // ---- write-out: the staged records are one contiguous, coalesced word copy ----
    tk_b200_candidate_t* out = k.cands + (size_t)frame * k.cap;
    if (frame_pre < k.cap) {
        const uint32_t nrec = min(min(local_tot, kDecStage), k.cap - frame_pre);
        TK_DEV_ASSERT(frame_pre + nrec <= k.cap && nrec <= kDecStage);
        uint32_t* dst = reinterpret_cast<uint32_t*>(out + frame_pre);
        for (uint32_t i = tid; i < nrec * 7; i += kDecThreads) dst[i] = stage[i];
    }
    // ---- overflow (more than kDecStage survivors in this CTA's slice): straight to global ----
    if (local_tot > kDecStage) {
        for (uint32_t w = warp; w < k.nwords; w += kDecWarps) {
            const uint32_t word = flags[w];
            uint32_t bits = word;
            while (bits) {
                const uint32_t b = __ffs(bits) - 1;
                bits &= bits - 1;
                const uint32_t lpos = wpre[w] + __popc(word & ((1u << b) - 1u));
                if (lpos < kDecStage) continue;
                const uint32_t pos = frame_pre + lpos;
                if (pos >= k.cap) { bits = 0; break; }             // later rows only have larger pos
                const uint32_t r = row0 + w * 32 + b;
                RowRegs q;
                load_row(k, fl + (size_t)r * k.L, lane, q);
                const tk_b200_candidate_t c = decode_row<ACT>(k, fl + (size_t)r * k.L, r, lane, my_e, q);
                if (lane == 0) out[pos] = c;
            }
        }
    }
#ifdef TK_B200_TRACE_POST
    if (k.trace.buf) __syncthreads();                              // the end stamp covers every thread's writes
#endif
    if (tr) { tstamp[5] = gtimer(); trace_put(k.trace, 3u, blockIdx.x, local_tot, tstamp, 6); }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    if constexpr (NMS) {
        // Every CTA of the cluster has written its slice of the frame's list (rank 0 also the count): release them to
        // the first CTA, which goes on to the NMS; the others are done.
        __syncthreads();
        asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
        if (rank != 0) return;
        asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (min(frame_tot, k.cap) <= kNmsSmall) nms_frame<kDecThreads>(tail.n, frame, dsm, kNmsSmall);
        else nms_frame<kDecThreads>(tail.n, frame, tail.scratch + (size_t)frame * nms_smem_bytes(tail.scratch_pitch), tail.scratch_pitch);
    }
}

template <int ACT>
static tk_error_code_t launch_decode(const DecK& k, const NmsTail* tail, uint32_t frames, size_t smem, cudaStream_t stream) {
    auto kern = tail ? (k.obj ? decode_kernel<ACT, true, true> : decode_kernel<ACT, false, true>)
                     : (k.obj ? decode_kernel<ACT, true, false> : decode_kernel<ACT, false, false>);
    const NmsTail t = tail ? *tail : NmsTail{};
    if (smem > 48 * 1024)
        TK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), TK_ERROR_GPU_CUDA_ERROR);
    // (The largest shared-memory carve-out, which would spare the SMs a reconfiguration after the pre-process kernel, was
    // measured: decode_kernel 42.3 -> 59.4 us.  The kernel wants its L1.)
    cudaLaunchConfig_t cfg{};
    const int cl = tail ? kDecClusterNms : kDecCluster;
    cfg.gridDim = dim3(frames * cl);
    cfg.blockDim = dim3(kDecThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_chain_hint() ? 2 : 1;
    TK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, k, t), TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE);
    return launch_status(tail ? "decode_kernel+nms" : "decode_kernel");
}

static tk_error_code_t decode_launch_common(const tk_b200_decode_params_t* p, const float* d_objectness, const NmsTail* tail,
                                            cudaStream_t stream, const uint32_t* d_bitmap = nullptr) {
    if (!p || !p->d_logits || !p->d_counts || (!p->d_candidates && p->max_candidates)) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_frames == 0) return TK_SUCCESS;
    if (p->num_proposals == 0) return TK_ERROR_INVALID_DIMENSIONS;      // dims check of tk_object_detector.c:325-336
    if ((unsigned)p->activation > TK_B200_ACT_SOFTMAX) return TK_ERROR_INVALID_ARGUMENT;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    DecK k{};
    k.logits = p->d_logits; k.obj = d_objectness; k.cands = p->d_candidates; k.counts = p->d_counts;
    k.bitmap = d_bitmap; k.nwf = (p->num_proposals + 31) / 32;
    k.P = p->num_proposals; k.C = p->num_classes; k.L = 5 + p->num_classes;
    const uint32_t cl = tail ? kDecClusterNms : kDecCluster;
    uint32_t per = (k.P + cl - 1) / cl;
    k.rows_per_cta = (per + 31) & ~31u;
    k.nwords = k.rows_per_cta / 32;
    k.cap = p->max_candidates;
    k.thr = p->confidence_threshold;
    k.trace = trace_target();
    size_t smem = (size_t)k.nwords * 8 + (p->activation == TK_B200_ACT_SOFTMAX ? (size_t)kDecWarps * k.C * 4 : 0);
    if (tail && smem < nms_smem_bytes(kNmsSmall)) smem = nms_smem_bytes(kNmsSmall);   // the NMS phase overlays the decode buffers
    if (smem > (size_t)dev->max_smem_optin - 1024 - 8192) {
        tk_error_set_detail("decode: %u proposals x %u classes need %zu B of shared memory", k.P, k.C, smem);
        return TK_ERROR_NOT_IMPLEMENTED;
    }
    switch (p->activation) {
        case TK_B200_ACT_RAW:     return launch_decode<TK_B200_ACT_RAW>(k, tail, p->num_frames, smem, stream);
        case TK_B200_ACT_SIGMOID: return launch_decode<TK_B200_ACT_SIGMOID>(k, tail, p->num_frames, smem, stream);
        default:                  return launch_decode<TK_B200_ACT_SOFTMAX>(k, tail, p->num_frames, smem, stream);
    }
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_kernels_decode_detections(const tk_b200_decode_params_t* p, cudaStream_t stream) {
    return tkb::decode_detections_packed(p, nullptr, stream);
}

tk_error_code_t tkb::decode_detections_packed(const tk_b200_decode_params_t* p, const float* d_objectness, cudaStream_t stream) {
    return decode_launch_common(p, d_objectness, nullptr, stream);
}

tk_error_code_t tkb::decode_detections_bitmap(const tk_b200_decode_params_t* p, const uint32_t* d_bitmap, cudaStream_t stream) {
    return decode_launch_common(p, nullptr, nullptr, stream, d_bitmap);
}

// tk_kernels_postprocess_fused, layout 0: decode_kernel with the NMS tail.  `n` comes filled from post.cu
// (postprocess_fused_packed validates and builds it); scratch: [frames][nms_smem_bytes(pitch)] when the cap is above 512.
tk_error_code_t tkb::decode_nms_cluster(const tk_b200_decode_params_t* p, const float* d_objectness, const void* nms_k,
                                        void* scratch, uint32_t scratch_pitch, cudaStream_t stream) {
    NmsTail t{};
    t.n = *static_cast<const NmsK*>(nms_k);
    t.scratch = static_cast<uint8_t*>(scratch);
    t.scratch_pitch = scratch_pitch;
    return decode_launch_common(p, d_objectness, &t, stream);
}
