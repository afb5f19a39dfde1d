// nms_frame.cuh — class-wise greedy IoU NMS of ONE frame by one CTA; shared by nms_kernel (nms.cu) and the fused
// co-resident post-process kernel (post.cu).  Semantics and design notes: see nms.cu.
#pragma once
#include "tk_common.cuh"

namespace tkb {

static constexpr int kNmsThreads = 512;     // nms_kernel's CTA; the fused post-process kernel runs nms_frame with 256
static constexpr uint32_t kNmsMaxK = 4096;   // nms_kernel: the most candidates one CTA holds in shared memory (200 KB)

struct NmsK {
    const tk_b200_candidate_t* cands;
    const uint32_t* counts;
    tk_b200_candidate_t* dets;
    uint32_t* num_dets;
    uint32_t cap, max_det, Kp;      // Kp: power of two >= cap (>= 64)
    float thr;
    int agnostic;
    int restore;
    float sx, sy;                   // src/in scale of the optional coordinate restore
    uint32_t num_frames;
    uint32_t select;                // 0: every frame; 1: only frames of <= kNmsSmall candidates; 2: only the others
    TraceBuf trace;                 // debug: per-frame phase stamps (tk_b200_debug_trace), buf == nullptr when off
};

static constexpr uint32_t kNmsSmall = 512;   // the register path's limit: 25 KB of shared memory per CTA instead of 200 KB

// src/vision/src/object_analysis.rs:71-91 (i32 arithmetic wraps like release-mode Rust).
__device__ __forceinline__ float iou_rect(int4 a, int4 b) {
    int x_left = max(a.x, b.x), y_top = max(a.y, b.y);
    int x_right = min((int)((unsigned)a.x + (unsigned)a.z), (int)((unsigned)b.x + (unsigned)b.z));
    int y_bottom = min((int)((unsigned)a.y + (unsigned)a.w), (int)((unsigned)b.y + (unsigned)b.w));
    if (x_right < x_left || y_bottom < y_top) return 0.0f;
    float inter = __fmul_rn((float)(int)((unsigned)x_right - (unsigned)x_left),
                            (float)(int)((unsigned)y_bottom - (unsigned)y_top));
    float a1 = (float)(int)((unsigned)a.z * (unsigned)a.w);
    float a2 = (float)(int)((unsigned)b.z * (unsigned)b.w);
    float uni = __fsub_rn(__fadd_rn(a1, a2), inter);
    return uni > 0.0f ? __fdiv_rn(inter, uni) : 0.0f;
}

// Work area: keys u64, box int4, pos u32, cls u32, dead u8, keys2 u64, segmask u64 per entry + one kept bit per entry.
__host__ __device__ constexpr size_t nms_smem_bytes(uint32_t Kp) { return (size_t)Kp * (8 + 16 + 4 + 4 + 1 + 8 + 8) + Kp / 8; }

// Bitonic sort of Kp 64-bit keys in shared memory (ascending), optional u32 payload.
template <int THREADS, bool PAYLOAD>
__device__ __forceinline__ void bitonic_sort(unsigned long long* keys, uint32_t* pay, uint32_t Kp, uint32_t tid) {
    for (uint32_t kk = 2; kk <= Kp; kk <<= 1) {
        for (uint32_t j = kk >> 1; j > 0; j >>= 1) {
            for (uint32_t t = tid; t < (Kp >> 1); t += THREADS) {
                uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
                uint32_t l = i | j;
                unsigned long long a = keys[i], b = keys[l];
                bool up = (i & kk) == 0;
                if ((a > b) == up) {
                    keys[i] = b; keys[l] = a;
                    if (PAYLOAD) { uint32_t pa = pay[i]; pay[i] = pay[l]; pay[l] = pa; }
                }
            }
            __syncthreads();
        }
    }
}

// Bitonic sort of 512 (key, payload) pairs held one per thread: partner distances below
// 32 are warp shuffles, the 10 larger ones go through shared memory (two buffers used
// alternately: one CTA barrier per exchange).  35 of the 45 stages need no barrier.
template <bool PAYLOAD>
__device__ __forceinline__ void bitonic_sort_regs512(unsigned long long& key, uint32_t& pay, unsigned long long* xk0,
                                                     unsigned long long* xk1, uint32_t* xp0, uint32_t* xp1, uint32_t tid) {
    uint32_t flip = 0;
#pragma unroll
    for (uint32_t kk = 2; kk <= 512; kk <<= 1) {
#pragma unroll
        for (uint32_t j = kk >> 1; j > 0; j >>= 1) {
            unsigned long long ok; uint32_t op = 0;
            if (j >= 32) {
                unsigned long long* xk = flip ? xk1 : xk0;
                uint32_t* xp = flip ? xp1 : xp0;
                flip ^= 1u;
                xk[tid] = key;
                if (PAYLOAD) xp[tid] = pay;
                __syncthreads();
                ok = xk[tid ^ j];
                if (PAYLOAD) op = xp[tid ^ j];
            } else {
                ok = __shfl_xor_sync(0xffffffffu, key, j);
                if (PAYLOAD) op = __shfl_xor_sync(0xffffffffu, pay, j);
            }
            const bool take_min = ((tid & j) == 0) == ((tid & kk) == 0);
            const bool swap = take_min ? (ok < key) : (ok > key);
            if (swap) { key = ok; pay = op; }
        }
    }
}

// `work`: nms_smem_bytes(pitch) bytes, 16-byte aligned — shared memory, or (frames whose candidates do not fit the
// CTA's shared memory) a per-CTA scratch in global memory: every access below is an ordinary load / store / atomic
// ordered by CTA barriers, so the same code runs on either.  THREADS: the CTA size (512: nms_kernel, which also has
// the one-candidate-per-thread register sort; 256: the co-resident post-process kernel).
template <int THREADS>
__device__ __forceinline__ void nms_frame(const NmsK& k, const uint32_t frame, uint8_t* dsm, const uint32_t pitch) {
    constexpr int kNmsWarps = THREADS / 32;
    __shared__ unsigned long long diag[64];
    __shared__ unsigned long long blk_kept;
    __shared__ uint32_t warp_tot[kNmsWarps];
    __shared__ uint32_t seg_overflow;
    __shared__ unsigned long long tstamp[6];                       // debug trace: start, sorts, segment scan, pair masks, resolve, end

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#ifdef TK_B200_TRACE_POST
    const bool tr = k.trace.buf != nullptr && tid == 0;
#else
    constexpr bool tr = false;                                     // phase stamps: `make trace` builds only
#endif
    if (tr) tstamp[0] = gtimer();
    const tk_b200_candidate_t* in = k.cands + (size_t)frame * k.cap;
    tk_b200_candidate_t* out = k.dets + (size_t)frame * k.max_det;
    const uint32_t K = min(k.counts[frame], k.cap);
    if (K == 0 || k.max_det == 0) { if (tid == 0) k.num_dets[frame] = 0; return; }
    if ((k.select == 1 && K > kNmsSmall) || (k.select == 2 && K <= kNmsSmall)) return;    // the other launch's frame
    uint32_t Kp = 64;
    while (Kp < K) Kp <<= 1;                                       // sort only what this frame needs
    TK_DEV_ASSERT(Kp <= pitch && (reinterpret_cast<uintptr_t>(dsm) & 15u) == 0);

    unsigned long long* keys = reinterpret_cast<unsigned long long*>(dsm);          // [Kp] score keys
    int4* box = reinterpret_cast<int4*>(keys + pitch);                               // [Kp] score order
    uint32_t* pos = reinterpret_cast<uint32_t*>(box + pitch);                        // [Kp] position in the input list
    uint32_t* cls = pos + pitch;                                                     // [Kp] score order
    uint8_t* dead = reinterpret_cast<uint8_t*>(cls + pitch);                         // [Kp] blocked path
    unsigned long long* keys2 = reinterpret_cast<unsigned long long*>(dead + pitch); // [Kp] (class << 32 | score rank)
    unsigned long long* segmask = keys2 + pitch;                                     // [Kp] class order
    unsigned long long* keptbits = segmask + pitch;                                  // [Kp / 64] kept flags by score rank

    for (uint32_t i = tid; i < (Kp >> 6); i += THREADS) keptbits[i] = 0ull;
    if (tid == 0) seg_overflow = 0;
    bool reg_sort = false;
    if constexpr (THREADS == 512) reg_sort = Kp <= 512;
    if (reg_sort) {
        // ---- register path: one candidate per thread, both sorts mostly in shuffles ----
        unsigned long long key = ~0ull; uint32_t p0 = tid;
        if (tid < K) {
            uint32_t ck = tk_dm_float_order_key(in[tid].confidence);
            key = ((unsigned long long)(~ck) << 32) | (uint32_t)in[tid].index;
        }
        // ascending key == confidence desc, index asc.  Equal keys (same confidence and
        // same index) can only be distinct input positions holding identical proposals;
        // any order among them gives the same output.
        bitonic_sort_regs512<true>(key, p0, keys, keys2, pos, cls, tid);
        __syncthreads();                                           // exchange buffers (pos, cls) are free again
        unsigned long long k2 = ~0ull;
        if (tid < K) {
            const tk_b200_candidate_t& c = in[p0];
            box[tid] = make_int4(c.bbox.x, c.bbox.y, c.bbox.w, c.bbox.h);
            const uint32_t cc = k.agnostic ? 0u : c.class_id;
            k2 = ((unsigned long long)cc << 32) | tid;
        }
        const uint32_t my_cls = (uint32_t)(k2 >> 32);
        uint32_t none = 0;
        // Second sort groups each class into one contiguous segment, score order inside
        // (a class id of 0xffffffff sorts with the padding but keeps a rank < K, so it is
        // told apart below by t < K).
        bitonic_sort_regs512<false>(k2, none, keys, keys2, nullptr, nullptr, tid);
        __syncthreads();
        if (tid < K) { pos[tid] = p0; cls[tid] = my_cls; dead[tid] = 0; }
        keys2[tid] = k2;
        __syncthreads();
    } else {
    // ---- score keys -----------------------------------------------------------
    for (uint32_t i = tid; i < Kp; i += THREADS) {
        unsigned long long key = ~0ull;
        if (i < K) {
            uint32_t ck = tk_dm_float_order_key(in[i].confidence);
            key = ((unsigned long long)(~ck) << 32) | (uint32_t)in[i].index;
        }
        keys[i] = key; pos[i] = i;
    }
    __syncthreads();
    bitonic_sort<THREADS, true>(keys, pos, Kp, tid);

    // ---- stage in score order; class keys -----------------------------------------
    for (uint32_t i = tid; i < Kp; i += THREADS) {
        unsigned long long k2 = ~0ull;
        if (i < K) {
            const tk_b200_candidate_t& c = in[pos[i]];
            box[i] = make_int4(c.bbox.x, c.bbox.y, c.bbox.w, c.bbox.h);
            const uint32_t cc = k.agnostic ? 0u : c.class_id;
            cls[i] = cc; dead[i] = 0;
            k2 = ((unsigned long long)cc << 32) | i;
        }
        keys2[i] = k2;
    }
    __syncthreads();
    bitonic_sort<THREADS, false>(keys2, nullptr, Kp, tid);
    }

    if (tr) tstamp[1] = gtimer();
    // ---- A. per-candidate suppressor mask inside its class segment -----------------
    // Suppression never crosses classes.  For the candidate at class-order position t
    // with R same-class predecessors, bit r of segmask[t] <=> the predecessor of
    // segment rank r has IoU > thr with it.  Segments longer than 64 (crowds of one
    // class, class-agnostic mode) take the blocked path below instead.
    //
    // The (candidate, predecessor) pairs are dealt out evenly: a frame's time is set by its
    // longest class (60 candidates of one class = 59 dependent IoUs for the last of them when
    // every candidate walks its own predecessors), while the pair list of a 480-candidate frame is
    // ~14 pairs per thread.  wpre[t] = number of pairs before candidate t (exclusive prefix of R).
    uint32_t* wpre = reinterpret_cast<uint32_t*>(keys);            // [K + 1]; the score keys are no longer needed
    {
        uint32_t carry_start = 0, carry_sum = 0;
        for (uint32_t base = 0; base < K; base += THREADS) {
            const uint32_t t = base + tid;
            const bool valid = t < K;
            const uint32_t c = valid ? (uint32_t)(keys2[t] >> 32) : 0u;
            const bool head = valid && (t == 0 || (uint32_t)(keys2[t - 1] >> 32) != c);
            uint32_t start = head ? t : 0u;                        // inclusive max-scan: start of my segment
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, start, d);
                if ((int)lane >= d) start = max(start, o);
            }
            if (lane == 31) warp_tot[warp] = start;
            __syncthreads();
            uint32_t prev = carry_start, all = carry_start;
#pragma unroll
            for (int w = 0; w < kNmsWarps; ++w) { const uint32_t v = warp_tot[w]; if ((uint32_t)w < warp) prev = max(prev, v); all = max(all, v); }
            start = max(start, prev);
            const uint32_t R = valid ? t - start : 0u;
            if (R >= 64) seg_overflow = 1;                         // a segment longer than 64
            uint32_t inc = R;                                      // inclusive sum-scan of R
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
                if ((int)lane >= d) inc += o;
            }
            __syncthreads();                                       // warp_tot: everyone has read the segment starts
            if (lane == 31) warp_tot[warp] = inc;
            __syncthreads();
            uint32_t before = carry_sum, tot = 0;
#pragma unroll
            for (int w = 0; w < kNmsWarps; ++w) { const uint32_t v = warp_tot[w]; before += (uint32_t)w < warp ? v : 0u; tot += v; }
            if (valid) {
                wpre[t] = before + inc - R;
                segmask[t] = 0ull;
                dead[t] = (uint8_t)(R == 0);                       // (fast path only) 1 = head of a segment
            }
            carry_start = all; carry_sum += tot;
            __syncthreads();
        }
        if (tid == 0) wpre[K] = carry_sum;
    }
    __syncthreads();

    uint32_t run = 0;
    if (tr) tstamp[2] = gtimer();
    if (!seg_overflow) {
        const uint32_t total = wpre[K];
        const uint32_t per = (total + THREADS - 1) / THREADS;
        uint32_t pi = tid * per;
        const uint32_t pend = min(pi + per, total);
        if (pi < pend) {
            uint32_t lo = 0, hi = K - 1;                           // last t with wpre[t] <= pi: the candidate owning pair pi
            while (lo < hi) {
                const uint32_t mid = (lo + hi + 1) >> 1;
                if (wpre[mid] <= pi) lo = mid; else hi = mid - 1;
            }
            uint32_t t = lo, w0 = wpre[t], R = wpre[t + 1] - w0, r = pi - w0;
            TK_DEV_ASSERT(t < K && R <= t && r < R && (uint32_t)keys2[t] < K);
            int4 me = box[(uint32_t)keys2[t]];
            unsigned long long m = 0;
            for (; pi < pend; ++pi) {
                if (iou_rect(box[(uint32_t)keys2[t - R + r]], me) > k.thr) m |= 1ull << r;
                if (++r == R) {
                    if (m) atomicOr(&segmask[t], m);
                    m = 0; r = 0;
                    if (pi + 1 < pend) {
                        do { ++t; R = wpre[t + 1] - wpre[t]; } while (R == 0);     // pairs remain, so a later candidate has some
                        me = box[(uint32_t)keys2[t]];
                    }
                }
            }
            if (m) atomicOr(&segmask[t], m);
        }
        __syncthreads();
        if (tr) tstamp[3] = gtimer();
        // ---- B. each segment head resolves its segment sequentially (<= 64 steps).  Keys and masks
        //         come in groups of four independent loads, so the chain per element is the kept-mask
        //         update alone, not a shared-memory round trip ----
        for (uint32_t t = tid; t < K; t += THREADS) {
            if (!dead[t]) continue;
            const uint32_t c = (uint32_t)(keys2[t] >> 32);
            unsigned long long kept = 0;
            bool open = true;
            for (uint32_t r0 = 0; open; r0 += 4) {
                unsigned long long kq[4], sq[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const uint32_t e = t + r0 + u;
                    const bool in = e < K;
                    kq[u] = in ? keys2[e] : 0ull;
                    sq[u] = in ? segmask[e] : 0ull;
                    if (!in) kq[u] = ~((unsigned long long)c << 32);   // a different class: ends the segment
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (!open) break;
                    if ((uint32_t)(kq[u] >> 32) != c) { open = false; break; }
                    if ((sq[u] & kept) == 0) {
                        kept |= 1ull << (r0 + u);
                        const uint32_t i = (uint32_t)kq[u];         // score rank
                        atomicOr(&keptbits[i >> 6], 1ull << (i & 63));
                    }
                }
            }
        }
        __syncthreads();
        if (tr) tstamp[4] = gtimer();
        // ---- C. kept bitmap -> ordered compaction in score order ---------------------
        for (uint32_t base = 0; base < K; base += THREADS) {
            const uint32_t i = base + tid;
            const bool keep = i < K && ((keptbits[i >> 6] >> (i & 63)) & 1ull);
            const uint32_t bal = __ballot_sync(0xffffffffu, keep);
            if (lane == 0) warp_tot[warp] = __popc(bal);
            __syncthreads();
            uint32_t before = 0, tot = 0;
#pragma unroll
            for (int w = 0; w < kNmsWarps; ++w) { uint32_t t = warp_tot[w]; before += (uint32_t)w < warp ? t : 0u; tot += t; }
            const uint32_t o = run + before + __popc(bal & ((1u << lane) - 1u));
            if (keep && o < k.max_det) {
                tk_b200_candidate_t c = in[pos[i]];
                if (k.restore) {
                    c.bbox.x = __float2int_rz(__fmul_rn((float)c.bbox.x, k.sx));
                    c.bbox.y = __float2int_rz(__fmul_rn((float)c.bbox.y, k.sy));
                    c.bbox.w = __float2int_rz(__fmul_rn((float)c.bbox.w, k.sx));
                    c.bbox.h = __float2int_rz(__fmul_rn((float)c.bbox.h, k.sy));
                }
                out[o] = c;
            }
            run += tot;
            __syncthreads();
            if (run >= k.max_det) break;
        }
        if (tid == 0) k.num_dets[frame] = min(run, k.max_det);
        if (tr) { tstamp[5] = gtimer(); trace_put(k.trace, 4u, frame, K, tstamp, 6); }
        return;
    }

    // =========================================================================
    // Blocked path (a class segment longer than 64): 64-column blocks of the
    // suppression matrix over the score-sorted list.
    // =========================================================================
    for (uint32_t i = tid; i < K; i += THREADS) dead[i] = 0;   // box/cls are staged in score order already
    __syncthreads();

    for (uint32_t b0 = 0; b0 < K; b0 += 64) {
        const uint32_t bn = min(64u, K - b0);
        // (a) diagonal block of the bitmask matrix: bit j of diag[i] <=> j < i suppresses i
        if (tid < 64) {
            unsigned long long m = 0;
            if (tid < bn && !dead[b0 + tid]) {
                const int4 me = box[b0 + tid]; const uint32_t mc = cls[b0 + tid];
                for (uint32_t j = 0; j < tid; ++j)
                    if (cls[b0 + j] == mc && !dead[b0 + j] && iou_rect(box[b0 + j], me) > k.thr) m |= 1ull << j;
            }
            diag[tid] = m;
        }
        __syncthreads();
        // (b) sequential resolve of the 64 candidates with the kept-mask in a register
        if (tid == 0) {
            unsigned long long kept = 0;
#pragma unroll 8
            for (uint32_t i = 0; i < 64; ++i) {
                bool alive = i < bn && !dead[b0 + i];
                if (alive && (diag[i] & kept) == 0) kept |= 1ull << i;
                else if (i < bn) dead[b0 + i] = 1;
            }
            blk_kept = kept;
        }
        __syncthreads();
        // (c) sweep: kept boxes of this block suppress every later candidate
        const unsigned long long kept = blk_kept;
        if (kept) {
            for (uint32_t i = b0 + 64 + tid; i < K; i += THREADS) {
                if (dead[i]) continue;
                const int4 me = box[i]; const uint32_t mc = cls[i];
                unsigned long long rem = kept;
                while (rem) {
                    const uint32_t j = __ffsll((long long)rem) - 1;
                    rem &= rem - 1;
                    if (cls[b0 + j] == mc && iou_rect(box[b0 + j], me) > k.thr) { dead[i] = 1; break; }
                }
            }
        }
        __syncthreads();
    }

    // ---- ordered compaction of the survivors --------------------------------
    for (uint32_t base = 0; base < K; base += THREADS) {
        const uint32_t i = base + tid;
        const bool keep = i < K && !dead[i];
        const uint32_t bal = __ballot_sync(0xffffffffu, keep);
        if (lane == 0) warp_tot[warp] = __popc(bal);
        __syncthreads();
        uint32_t before = 0, tot = 0;
#pragma unroll
        for (int w = 0; w < kNmsWarps; ++w) { uint32_t t = warp_tot[w]; before += (uint32_t)w < warp ? t : 0u; tot += t; }
        const uint32_t o = run + before + __popc(bal & ((1u << lane) - 1u));
        if (keep && o < k.max_det) {
            tk_b200_candidate_t c = in[pos[i]];
            if (k.restore) {
                c.bbox.x = __float2int_rz(__fmul_rn((float)c.bbox.x, k.sx));
                c.bbox.y = __float2int_rz(__fmul_rn((float)c.bbox.y, k.sy));
                c.bbox.w = __float2int_rz(__fmul_rn((float)c.bbox.w, k.sx));
                c.bbox.h = __float2int_rz(__fmul_rn((float)c.bbox.h, k.sy));
            }
            out[o] = c;
        }
        run += tot;
        __syncthreads();
        if (run >= k.max_det) break;
    }
    if (tid == 0) k.num_dets[frame] = min(run, k.max_det);
}


}  // namespace tkb
