// preprocess.cu — fused frame ingest -> bilinear resize -> normalise -> NCHW pack.
//
// Replaces  tk_preprocessor_resize_and_normalize_to_chw
//             (src/vision/tk_image_preprocessor.c:21-166, the parity target) and
//           tk_cuda_preprocess_image_kernel / tk_kernels_preprocess_image
//             (src/gpu/cuda/tk_cuda_kernels.cu:102-169, :252-303).
//
// B200 design (HBM-bound byte work, no tensor cores):
//   * The global output-row space [0, N*H_out) is cut into one contiguous run
//     per CTA (grid = resident CTAs), so every source row is fetched once per
//     CTA and the load is balanced to +-1 row.
//   * Source rows stream through a shared-memory ring filled by TMA 1-D bulk
//     copies (cp.async.bulk + mbarrier complete_tx).  The CTA is warp-specialised:
//     ONE PRODUCER WARP walks the CTA's source-row sequence and issues the copies
//     as slots free up; the consumer warps wait on a slot's `full` barrier, read,
//     and arrive on its `empty` barrier.  No CTA-wide barrier after set-up, so the
//     warps drift apart and loads, gathers and stores overlap.  An unaligned
//     source (TMA's 16-byte rules) is copied by the producer warp with aligned
//     32-bit loads + funnel shift into the same ring.
//   * y_ratio >= 2 (4K): only the two source rows each output row samples are
//     staged ("sparse"); rows nothing samples are never fetched.
//   * Each thread produces 4 consecutive output pixels x 3 channels from the
//     ring (3 LDS.32 + funnel shift per source row and pixel, bank-conflict-free
//     at the 1080p->640 ratio) and writes one 128-bit (fp32) or 64-bit (fp16)
//     streaming store per channel plane.
//   * EXACT arithmetic reproduces the reference's fp32 operation order with no
//     contraction (bit-identical to its AVX build); the divide by std is the
//     Markstein fma sequence on a host-computed correctly rounded reciprocal.
#include "decode_row.cuh"

#include <atomic>
#include <stdio.h>
#include <stdlib.h>

namespace tkb {

enum { FMT_RGB8 = 0, FMT_RGBA8 = 1, FMT_I420 = 2 };

// Decode's pass 1 riding in this kernel (tk_b200_pipeline_process_device): one extra warp per CTA thresholds the
// objectness word of a contiguous slice of the batch's logit rows (ld.global.nc.L2::64B, 8 loads in flight per lane)
// and writes the warp-ballot words to `bitmap` ([frames][nwf] u32, bit b of word w of a frame = row 32 w + b passes),
// which decode_kernel then loads instead of scanning.  That scan is the HBM half of decode — 1.6 M scattered 64-byte
// fetches, 23 us of the 41 us decode_kernel takes behind this kernel — and it needs nothing this kernel's SMs are short
// of: 64 more threads per SM, no shared memory, loads that wait on DRAM while the consumer warps issue arithmetic.
struct ScanK {
    const float* logits;      // nullptr: no scanner warp
    uint32_t* bitmap;
    uint32_t P, L, nwf;       // proposals per frame, row length, bitmap words per frame = ceil(P / 32)
    uint32_t total_words;     // frames * nwf
    float thr;
    int act;                  // tk_b200_activation_e: RAW thresholds the word itself, the others its sigmoid
};

struct PreK {
    const uint8_t* in_base;
    const uint8_t* const* in_ptrs;
    unsigned long long frame_pitch;
    void* out;
    uint32_t num_frames, src_w, src_h, src_stride, dst_w, dst_h;
    uint32_t new_w, new_h, pad_left, pad_top;     // letterbox region inside dst (== dst for stretch)
    float x_ratio, y_ratio;
    float mean[3], stdv[3], rstd[3], fscale[3], fbias[3];
    float pad_value;
    uint32_t h_scale_rg, h_bias_rg, h_scale_bb, h_bias_bb;   // HALF arithmetic: half2 {R,G} and {B,B} of 1/(255 std), -mean/std
    uint32_t s_log2;        // ring slots = 1 << s_log2
    uint32_t slot_bytes;    // bytes per ring slot (multiple of 16)
    uint32_t copy_bytes;    // bytes per bulk copy of one (luma) row
    uint32_t chroma_copy_bytes, chroma_pitch, chroma_h;
    unsigned long long u_off, v_off;   // I420 plane offsets from the frame base
    uint32_t u_slot_off, v_slot_off;   // I420 offsets inside a ring slot
    uint32_t tab_w;         // dst_w rounded up to 4
    uint32_t ring_off;      // byte offset of the ring in dynamic smem
    uint32_t tx, ty;        // consumer threads: tx along the row (multiple of 32), ty row groups in flight
    uint32_t pdl;           // 1: launch as a programmatic dependent (pipeline only, see launch_variant)
    uint32_t trigger;       // 1: the pipeline's decode_kernel follows on the stream and may be launched under this grid's tail
    uint32_t vec_store;     // 1: dst_w % 4 == 0 and the output base is 16-byte aligned
    uint32_t sparse;        // 1: y_ratio >= 2 — only the two source rows of each output row are staged
    uint32_t swap_rb;       // 1: BGR byte order — source channel c lands in tensor plane 2 - c
    float one;              // 1.0f at run time (see sum2)
    unsigned long long cplane[3];   // byte offset of the tensor plane that SOURCE channel c is written to (BGR: 2 - c)
    TraceBuf trace;         // debug: per-CTA record (tk_b200_debug_trace), buf == nullptr when off
    ScanK scan;
    // Work distribution: global output rows [0, static_rows) are cut into one contiguous run per CTA; the rest,
    // [static_rows, N * dst_h), is handed out in chunks of dyn_rows rows from an atomic counter (g_pre_work[dyn_slot]) to
    // whichever CTA is through its run — CTAs of one launch differ by up to 45 % in speed (profiles/r2_experiments.md §15).
    unsigned long long static_rows;
    uint32_t dyn_rows, dyn_chunks, dyn_slot;
};

// Chunk counters of the launches in flight (a launch takes the next slot; the last CTA to finish clears its slot's two
// words, so every launch finds zeros) — file-scope device memory: no workspace in the C-ABI, no memset on the stream.
static constexpr uint32_t kPreSlots = 64;
static __device__ unsigned int g_pre_work[kPreSlots];
static __device__ unsigned int g_pre_done[kPreSlots];

static constexpr float kBias = 8388608.0f;   // 2^23: 0x4B0000bb == 2^23 + bb

__device__ __forceinline__ float magic_byte(uint32_t word, uint32_t sel) {
    // PRMT: {byte sel of word, 0, 0, 0x4B} -> float 2^23 + byte
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440u | sel));
}

// Column constants of one output pixel, worked out once per column (load_cols in the kernel) instead of once per pixel:
//   RGB8   off = byte offset of the 4-byte word holding the left source pixel's first byte, aux = its bit position in that word
//   RGBA8  off = byte offset of the left source pixel
//   I420   off = luma index of the left source pixel, aux = chroma index of the left pixel, bit 31: the right pixel's is one more
template <int FMT>
__device__ __forceinline__ void col_consts(const PreK& p, int xb, int& off, uint32_t& aux) {
    if constexpr (FMT == FMT_RGB8) { off = xb & ~3; aux = ((uint32_t)xb & 3u) << 3; }
    else if constexpr (FMT == FMT_RGBA8) { off = xb; aux = 0; }
    else {
        const int cw = (int)(p.src_w >> 1);
        off = xb;
        const int c0 = min(xb >> 1, cw - 1), c1 = min((xb + 1) >> 1, cw - 1);
        aux = (uint32_t)c0 | ((uint32_t)(c1 - c0) << 31);                 // c1 - c0 is 0 or 1
    }
}

// Fetch the 2x2 source neighbourhood of one output pixel as biased floats
// m[k][c] = 2^23 + value, k = {top-left, top-right, bottom-left, bottom-right}.
// XU (RGB8, EXACT): the bytes per source row that sit in byte 2 / byte 3 of a word after the funnel shifts — B of the
// left pixel, R, G, B of the right one — are converted by the conversion unit (I2F.U8 with a byte selector, no ALU-pipe
// instruction) and arrive as PLAIN floats p, not 2^23 + p: the kernel is bound by what it issues to the ALU pipe
// (profiles/r2_experiments.md §9), the conversion unit is idle.  blend_norm2 multiplies those taps directly.  Four of
// the six bytes is the balance: 94.8 -> 92.5 us with two, 90.4 us with four, 106 us with all six (the unit saturates).
template <int FMT, bool XU = false>
__device__ __forceinline__ void fetch_quad(const PreK& p, const uint8_t* row0, const uint8_t* row1,
                                           int off, uint32_t aux, float (&m)[4][3]) {
    if constexpr (FMT == FMT_RGB8) {
        const uint32_t* a = reinterpret_cast<const uint32_t*>(row0 + off);
        const uint32_t* b = reinterpret_cast<const uint32_t*>(row1 + off);
        uint32_t a0 = a[0], a1 = a[1], a2 = a[2];
        uint32_t b0 = b[0], b1 = b[1], b2 = b[2];
        const uint32_t sh = aux;                         // 0, 8, 16, 24
        uint32_t alo = __funnelshift_r(a0, a1, sh), ahi = __funnelshift_r(a1, a2, sh);
        uint32_t blo = __funnelshift_r(b0, b1, sh), bhi = __funnelshift_r(b1, b2, sh);
        m[0][0] = magic_byte(alo, 0); m[0][1] = magic_byte(alo, 1);
        m[2][0] = magic_byte(blo, 0); m[2][1] = magic_byte(blo, 1);
        if constexpr (XU) {
            // window bytes 2..5 in one more word: G and B of the right pixel land in its bytes 2 and 3 as well
            const uint32_t amid = __funnelshift_r(alo, ahi, 16), bmid = __funnelshift_r(blo, bhi, 16);
            m[0][2] = (float)((alo >> 16) & 0xffu); m[1][0] = (float)(alo >> 24);
            m[1][1] = (float)((amid >> 16) & 0xffu); m[1][2] = (float)(amid >> 24);
            m[2][2] = (float)((blo >> 16) & 0xffu); m[3][0] = (float)(blo >> 24);
            m[3][1] = (float)((bmid >> 16) & 0xffu); m[3][2] = (float)(bmid >> 24);
        } else {
            m[0][2] = magic_byte(alo, 2); m[1][0] = magic_byte(alo, 3);
            m[1][1] = magic_byte(ahi, 0); m[1][2] = magic_byte(ahi, 1);
            m[2][2] = magic_byte(blo, 2); m[3][0] = magic_byte(blo, 3);
            m[3][1] = magic_byte(bhi, 0); m[3][2] = magic_byte(bhi, 1);
        }
    } else if constexpr (FMT == FMT_RGBA8) {
        const uint32_t* a = reinterpret_cast<const uint32_t*>(row0 + off);
        const uint32_t* b = reinterpret_cast<const uint32_t*>(row1 + off);
        uint32_t a0 = a[0], a1 = a[1], b0 = b[0], b1 = b[1];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            m[0][c] = magic_byte(a0, c); m[1][c] = magic_byte(a1, c);
            m[2][c] = magic_byte(b0, c); m[3][c] = magic_byte(b1, c);
        }
    } else {
        // I420: off is the luma pixel index.  BT.601 full range, 16.16 fixed point,
        // nearest chroma — the integer arithmetic of oracle_i420_to_rgb8.
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint8_t* row = (k & 2) ? row1 : row0;
            const int x = off + (k & 1);
            const int cx = (int)((aux & 0x7fffffffu) + ((k & 1) ? (aux >> 31) : 0u));
            // (yy*65536 + K*(c-128) + 32768) >> 16 == yy + ((K*c + 32768 - 128*K) >> 16): the luma term is a
            // multiple of 65536 and >> is a floor.  add + clamp to [0, 255] is one VIADDMNMX.RELU (DPX).
            const int yy = row[x];
            const int u = row[p.u_slot_off + cx], v = row[p.v_slot_off + cx];
            const int dr = (91881 * v + (32768 - 128 * 91881)) >> 16;
            const int dg = (-22554 * u - 46802 * v + (32768 + 128 * (22554 + 46802))) >> 16;
            const int db = (116130 * u + (32768 - 128 * 116130)) >> 16;
            const int r = __viaddmin_s32_relu(yy, dr, 255);
            const int g = __viaddmin_s32_relu(yy, dg, 255);
            const int b = __viaddmin_s32_relu(yy, db, 255);
            m[k][0] = __uint_as_float(0x4B000000u | (uint32_t)r);
            m[k][1] = __uint_as_float(0x4B000000u | (uint32_t)g);
            m[k][2] = __uint_as_float(0x4B000000u | (uint32_t)b);
        }
    }
}

// Reference arithmetic, src/vision/tk_image_preprocessor.c:63-66 then :84-98.
template <bool EXACT>
__device__ __forceinline__ float blend_norm(const PreK& p, int c, float m1, float m2, float m3, float m4,
                                            float xd, float omx, float yd, float omy) {
    if constexpr (EXACT) {
        float p1 = __fsub_rn(m1, kBias), p2 = __fsub_rn(m2, kBias);
        float p3 = __fsub_rn(m3, kBias), p4 = __fsub_rn(m4, kBias);
        float t1 = __fmul_rn(__fmul_rn(p1, omx), omy);
        float t2 = __fmul_rn(__fmul_rn(p2, xd), omy);
        float t3 = __fmul_rn(__fmul_rn(p3, omx), yd);
        float t4 = __fmul_rn(__fmul_rn(p4, xd), yd);
        float v = __fadd_rn(__fadd_rn(__fadd_rn(t1, t2), t3), t4);
        float t = __fsub_rn(__fmul_rn(v, 1.0f / 255.0f), p.mean[c]);
        // correctly rounded t / std: q = t*y, r = t - q*std (exact), q' = q + r*y
        float q = __fmul_rn(t, p.rstd[c]);
        float r = __fmaf_rn(-q, p.stdv[c], t);
        return __fmaf_rn(r, p.rstd[c], q);
    } else {
        float top = __fmaf_rn(xd, m2 - m1, m1 - kBias);
        float bot = __fmaf_rn(xd, m4 - m3, m3 - kBias);
        float v = __fmaf_rn(yd, bot - top, top);
        return __fmaf_rn(v, p.fscale[c], p.fbias[c]);
    }
}

// The same arithmetic for one channel of TWO neighbouring output pixels at once, in packed fp32
// (FADD2 / FMUL2 / FFMA2, sm_100: two IEEE round-to-nearest lanes per issued instruction) — every
// lane performs exactly the scalar sequence above, so the results are bit-identical to it.
// The bytes arrive as m = 2^23 + p (magic_byte).  The reference's first products p*(1-xd), p*xd are formed as
// fma(m, w, -2^23*w): -2^23*w is exact (a power-of-two scaling), so the fma computes m*w - 2^23*w = p*w exactly and
// rounds ONCE — the same value as fmul(p, w) — and the four subtractions that would turn m back into p disappear.
// cxd = -2^23*xd, comx = -2^23*(1-xd).
// plain_l / plain_r: the left (m1, m3) / right (m2, m4) taps are plain byte values (fetch_quad<.., XU>): p * w directly.
template <bool EXACT>
__device__ __forceinline__ float2 blend_norm2(const PreK& p, int c, float2 m1, float2 m2, float2 m3, float2 m4,
                                              float2 xd, float2 omx, float2 cxd, float2 comx, float yd, float omy,
                                              bool plain_l = false, bool plain_r = false) {
    if constexpr (EXACT) {
        const float2 yd2 = f2s(yd), omy2 = f2s(omy);
        const float2 t1 = mul2(plain_l ? mul2(m1, omx) : fma2(m1, omx, comx), omy2);
        const float2 t2 = mul2(plain_r ? mul2(m2, xd) : fma2(m2, xd, cxd), omy2);
        const float2 t3 = mul2(plain_l ? mul2(m3, omx) : fma2(m3, omx, comx), yd2);
        const float2 t4 = mul2(plain_r ? mul2(m4, xd) : fma2(m4, xd, cxd), yd2);
        const float2 v = sum2(sum2(sum2(t1, t2, p.one), t3, p.one), t4, p.one);
        const float2 t = sum2(mul2(v, f2s(1.0f / 255.0f)), f2s(-p.mean[c]), p.one);        // x - mean == x + (-mean)
        const float2 q = mul2(t, f2s(p.rstd[c]));
        const float2 r = fma2(q, f2s(-p.stdv[c]), t);                                  // -q*std == q*(-std)
        return fma2(r, f2s(p.rstd[c]), q);
    } else {
        const float2 nb = f2s(-kBias);
        const float2 top = fma2(xd, add2(m2, f2(-m1.x, -m1.y)), add2(m1, nb));
        const float2 bot = fma2(xd, add2(m4, f2(-m3.x, -m3.y)), add2(m3, nb));
        const float2 v = fma2(f2s(yd), add2(bot, f2(-top.x, -top.y)), top);
        return fma2(v, f2s(p.fscale[c]), f2s(p.fbias[c]));
    }
}

template <bool EXACT>
__device__ __forceinline__ float norm_only(const PreK& p, int c, float v) {
    if constexpr (EXACT) {
        float t = __fsub_rn(__fmul_rn(v, 1.0f / 255.0f), p.mean[c]);
        float q = __fmul_rn(t, p.rstd[c]);
        float r = __fmaf_rn(-q, p.stdv[c], t);
        return __fmaf_rn(r, p.rstd[c], q);
    } else {
        return __fmaf_rn(v, p.fscale[c], p.fbias[c]);
    }
}

// Source row of output row h (src/vision/tk_image_preprocessor.c:49,51), clamped so row+1 exists.
__device__ __forceinline__ int src_row_of(const PreK& p, uint32_t h) {
    return min(__float2int_rz(__fmul_rn(p.y_ratio, (float)(h - p.pad_top))), (int)p.src_h - 2);
}

// Inside-region output rows [a, b) of an item covering output rows [h0, h1), and the
// ring sequence the item occupies:
//   dense  (y_ratio < 2): every source row ylo .. y(b-1)+1 once, seq = row - ylo;
//   sparse (y_ratio >= 2): rows y(h), y(h)+1 per output row, seq = 2*(h - a) + {0,1} —
//          source rows no output row samples are never fetched.
__device__ __forceinline__ void item_src_rows(const PreK& p, uint32_t h0, uint32_t h1, int& ylo, int& nrows, uint32_t& a_out) {
    uint32_t a = max(h0, p.pad_top), b = min(h1, p.pad_top + p.new_h);
    a_out = a;
    if (a >= b) { ylo = 0; nrows = 0; return; }
    ylo = src_row_of(p, a);
    nrows = p.sparse ? (int)(2u * (b - a)) : src_row_of(p, b - 1) + 2 - ylo;
}

// Copy `nb` bytes from a possibly unaligned global address into 4-byte-aligned shared
// memory, one warp, aligned 32-bit loads + funnel shift (fallback when TMA's 16-byte
// rules do not hold).  Never reads a global word that holds no byte of [src, src+nb).
__device__ __forceinline__ void warp_copy_unaligned(uint8_t* dst, const uint8_t* src, uint32_t nb, uint32_t lane) {
    const uint32_t shift = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 3u);
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(src - shift);
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst);
    const uint32_t nd = (nb + 3) >> 2;
    for (uint32_t i = lane; i < nd; i += 32) {
        uint32_t lo = __ldg(sw + i);
        uint32_t hi = (shift && 4u * (i + 1) < shift + nb) ? __ldg(sw + i + 1) : 0u;
        dw[i] = __funnelshift_r(lo, hi, shift * 8);
    }
}

// Block = consumer threads (tx * ty, tx a multiple of 32) + one producer warp.
//   consumers: warp (cy, *) produces output rows h0+cy, h0+cy+ty, ...; before a row it
//              publishes `progress[warp]` = the first ring sequence number the row
//              needs (everything below is released), waits for its two source rows on
//              their `full` barriers and computes.
//   producer : walks the CTA's source-row sequence; sequence q may overwrite slot
//              q mod S once q - S < min(progress[*]); lane 0 then issues the TMA bulk
//              copy (or the warp copies) that completes full[slot].
// There is no CTA-wide barrier after set-up.
// HALF arithmetic (fp16 output, RGB8 / RGBA8): the 2x2 neighbourhood as half2 {R,G} and {B,B}
// pairs, 0x6400 | byte == 1024 + byte exactly, differences exact, lerps and the
// normalisation in HFMA2 — half the math instructions of the fp32 lerp.  Error bound: DESIGN.md §3.
template <int FMT>
__device__ __forceinline__ void fetch_pairs_half(const uint8_t* row, int off, uint32_t aux, uint32_t& l_rg, uint32_t& l_bb,
                                                 uint32_t& r_rg, uint32_t& r_bb) {
    const uint32_t C64 = 0x64646464u;
    uint32_t lw, rw;                                   // words starting at the left / right pixel
    if constexpr (FMT == FMT_RGB8) {
        const uint32_t* a = reinterpret_cast<const uint32_t*>(row + off);
        const uint32_t a0 = a[0], a1 = a[1], a2 = a[2];
        const uint32_t sh = aux;
        lw = __funnelshift_r(a0, a1, sh);
        rw = __funnelshift_r(lw, __funnelshift_r(a1, a2, sh), 24);
    } else {
        const uint32_t* a = reinterpret_cast<const uint32_t*>(row + off);
        lw = a[0]; rw = a[1];
    }
    l_rg = __byte_perm(lw, C64, 0x4140); l_bb = __byte_perm(lw, C64, 0x4242);
    r_rg = __byte_perm(rw, C64, 0x4140); r_bb = __byte_perm(rw, C64, 0x4242);
}
__device__ __forceinline__ __half2 as_h2(uint32_t u) { return *reinterpret_cast<__half2*>(&u); }
__device__ __forceinline__ uint32_t h2_bits(__half2 h) { return *reinterpret_cast<uint32_t*>(&h); }
__device__ __forceinline__ __half2 lerp_pair_half(uint32_t tl, uint32_t tr, uint32_t bl, uint32_t br, __half2 xd2, __half2 yd2) {
    // 1152 = 1024 + 128: the lerp runs on CENTRED values (byte - 128, |v| <= 128), which halves the fp16 rounding step of
    // every intermediate against values up to 255; the +128 is folded into the bias constant on the host.
    const __half2 k1152 = as_h2(0x64806480u);
    const __half2 top = __hfma2(xd2, __hsub2(as_h2(tr), as_h2(tl)), __hsub2(as_h2(tl), k1152));
    const __half2 bot = __hfma2(xd2, __hsub2(as_h2(br), as_h2(bl)), __hsub2(as_h2(bl), k1152));
    return __hfma2(yd2, __hsub2(bot, top), top);
}

enum { AR_EXACT = 0, AR_FAST = 1, AR_HALF = 2 };
#ifndef TK_PRE_XU
#define TK_PRE_XU 1
#endif
static constexpr bool kPreXu = TK_PRE_XU != 0;   // byte -> float of four taps per pixel on the conversion unit (fetch_quad)

#ifdef TK_B200_CHECKED
// The ring slot the consumer is about to read / has just read must hold exactly source row `row` of the frame.
template <int FMT>
__device__ __noinline__ void check_ring_row(const PreK& p, const uint8_t* base, int row, const uint8_t* slot, uint32_t lane) {
    TK_DEV_ASSERT(row >= 0 && row < (int)p.src_h);
    const uint32_t nb = (FMT == FMT_RGB8) ? p.src_w * 3 : (FMT == FMT_RGBA8 ? p.src_w * 4 : p.src_w);
    const volatile uint8_t* g = base + (size_t)row * p.src_stride;
    for (uint32_t i = lane; i < nb; i += 32) TK_DEV_ASSERT(slot[i] == g[i]);
    if constexpr (FMT == FMT_I420) {
        const uint32_t cr = min((uint32_t)row >> 1, p.chroma_h - 1);
        const volatile uint8_t* gu = base + p.u_off + (size_t)cr * p.chroma_pitch;
        const volatile uint8_t* gv = base + p.v_off + (size_t)cr * p.chroma_pitch;
        for (uint32_t i = lane; i < (p.src_w >> 1); i += 32) {
            TK_DEV_ASSERT(slot[p.u_slot_off + i] == gu[i]);
            TK_DEV_ASSERT(slot[p.v_slot_off + i] == gv[i]);
        }
    }
}
#endif

// Row records: everything a consumer warp needs to know about one output row, worked out ONCE by the producer warp
// (one lane per row, up to 32 rows per pass) instead of by each of the row's five consumer warps — source-row index,
// y weight, ring slots, barrier parities, output address.  ~95 instructions of per-row set-up in every consumer warp
// become two 16-byte shared-memory loads (profiles/r2_experiments.md §9: the EXACT kernel is bound by the instructions
// it issues, ALU-pipe ones first).
//   word 0, 1  byte offsets (from the start of dynamic shared memory) of the ring slots holding source rows y, y + 1
//   word 2     y weight (fp32 bits)
//   word 3     bit 0: the row lies inside the resized region (else it is letterbox padding: no source rows);
//              bit 1, 2: phase parities of the two slots' `full` barriers for these rows
//   word 4, 5  shared-window addresses of those two barriers
//   word 6, 7  address of the output row in tensor plane 0 of its frame
// Record n (the CTA's n-th output row) lives in entry n mod kRecs and is published through `rec_ready` once both its
// source rows have been armed; entry n is reused for row n + kRecs only when every consumer warp has moved past row n.
static constexpr uint32_t kRecs = 32;
static constexpr uint32_t kSmemRecs = 512;                        // 32 x 32 B
static constexpr uint32_t kSmemQ0 = kSmemRecs + kRecs * 32;       // [kRecs] u32: first ring sequence number row n needs (producer only)
[[maybe_unused]] static constexpr uint32_t kSmemChk = kSmemQ0 + kRecs * 4;   // CHECKED build: [kRecs] {frame, source row, n}
static constexpr uint32_t kSmemRq = 2048;                         // [8] {first, end} global output rows of the work ranges in flight (producer only)
static constexpr uint32_t kSmemTabs = 2176;                       // column tables

// Block: tx * ty consumer threads (320 at a 640-pixel row) + the producer warp (+ the scanner warp, see ScanK).
// (512-thread CTAs — three row groups at <= 64 registers, 32 resident warps per SM instead of 22 — were built and
// measured: 112.5 us against 112.0 us for 64 x 1080p EXACT -> fp16; the kernel is not short of warps.  Two output rows
// per warp iteration: 117.3 us against 111.1 us — the set-up it meant to share is per row, see the records above.)
template <int FMT, bool F16, int AR, bool LBOX, bool TMA>
__global__ void __launch_bounds__(384, 2) preprocess_kernel(const PreK p) {
    constexpr bool EXACT = (AR == AR_EXACT);
    static_assert(AR != AR_HALF || (F16 && FMT != FMT_I420), "HALF arithmetic: fp16 output, RGB8/RGBA8 input");
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);                 // 32 x 8 B
    volatile uint32_t* progress = reinterpret_cast<volatile uint32_t*>(full + 32);   // per consumer warp (<= 32): the row it is on
    volatile uint32_t* rec_ready = progress + 32;                       // number of row records published so far
    uint4* recs = reinterpret_cast<uint4*>(smem + kSmemRecs);
    volatile uint32_t* q0s = reinterpret_cast<volatile uint32_t*>(smem + kSmemQ0);
    int* xb_tab = reinterpret_cast<int*>(smem + kSmemTabs);
    float* xd_tab = reinterpret_cast<float*>(smem + kSmemTabs + 4 * p.tab_w);
    uint8_t* ring = smem + p.ring_off;

    const uint32_t tid = threadIdx.x;
    const uint32_t ncons = p.tx * p.ty;
    const uint32_t S = 1u << p.s_log2, smask = S - 1;
    unsigned long long t_trace[2];
    if (p.trace.buf && tid == 0) t_trace[0] = gtimer();
    // PDL chain (opt-in): decode_kernel, which starts on griddepcontrol.wait, may be LAUNCHED once every CTA here is done
    // with its rows — its CTA launch ramp then runs under this grid's drain instead of after it.  (Announcing it at the
    // top of this kernel was measured: decode's resident-and-waiting CTAs take registers and slots from the CTAs still
    // working, 183 us against 165 us per step.)
    auto finish = [&]() {
        if (p.trigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        // Possibly launched early behind an unrelated kernel (see launch_variant): complete only after it has.
        asm volatile("griddepcontrol.wait;" ::: "memory");
    };

    if (tid == 0) {
        for (uint32_t s = 0; s < S; ++s) mbar_init(&full[s], 1);
        mbar_fence_init();
    }
    if (tid < 32) progress[tid] = tid < (ncons >> 5) ? 0u : 0xffffffffu;
    if (tid == 0) *rec_ready = 0u;
    // Column table: byte offset of the left source pixel and x weight
    // (src/vision/tk_image_preprocessor.c:48,50,52).
    for (uint32_t w = tid; w < p.tab_w; w += blockDim.x) {
        uint32_t wc = min(w, p.dst_w - 1);
        int xb = -1; float xd = 0.0f;
        if (!LBOX || (wc - p.pad_left) < p.new_w) {
            float gx = __fmul_rn(p.x_ratio, (float)(wc - p.pad_left));
            int x = min(__float2int_rz(gx), (int)p.src_w - 2);
            xd = __fsub_rn(gx, (float)x);
            xb = (FMT == FMT_RGB8) ? 3 * x : (FMT == FMT_RGBA8 ? 4 * x : x);
        }
        if constexpr (AR == AR_HALF) xd = __uint_as_float(h2_bits(__float2half2_rn(xd)));   // half2 {xd, xd} bit pattern
        xb_tab[w] = xb; xd_tab[w] = xd;
    }
    __syncthreads();

    // ======================= scanner warp (decode pass 1, see ScanK) =================
    if (tid >= ncons + 32) {
        const ScanK& s = p.scan;
        const uint32_t lane = tid & 31;
        const uint32_t w0 = (uint32_t)((unsigned long long)s.total_words * blockIdx.x / gridDim.x);
        const uint32_t w1 = (uint32_t)((unsigned long long)s.total_words * (blockIdx.x + 1) / gridDim.x);
        uint32_t frame = w0 / s.nwf, wi = w0 - frame * s.nwf;      // word w = frame * nwf + wi
        for (uint32_t wb = w0; wb < w1; wb += 8) {
            float obj[8];
            uint32_t f = frame, i = wi;
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint32_t r = i * 32 + lane;
                obj[u] = (wb + u < w1 && r < s.P) ? ld_obj(s.logits + ((size_t)f * s.P + r) * s.L + 4) : __int_as_float(0x7fc00000);
                if (++i == s.nwf) { i = 0; ++f; }
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const float a = s.act == TK_B200_ACT_RAW ? obj[u] : tk_dm_sigmoidf(obj[u]);
                const uint32_t b = __ballot_sync(0xffffffffu, a > s.thr);       // NaN never passes
                if (lane == (uint32_t)u && wb + u < w1) s.bitmap[wb + u] = b;
            }
            frame = f; wi = i;
        }
        finish();
        return;
    }

    // This CTA's contiguous run of the statically distributed output rows (the rest comes in chunks, see PreK).
    const unsigned long long total = (unsigned long long)p.num_frames * p.dst_h;
    const unsigned long long g0 = p.static_rows * blockIdx.x / gridDim.x;
    const unsigned long long g1 = p.static_rows * (blockIdx.x + 1) / gridDim.x;
    constexpr uint32_t ESZ = F16 ? 2 : 4;
    constexpr uint32_t kRecEnd = 8u;             // record flag: no more rows for this CTA

    // ======================= producer warp =======================================
    // Two cursors over the CTA's items (an item = the CTA's rows inside one frame): the ARMING cursor walks the source-row
    // sequence and issues the copies as slots free up; the RECORD cursor walks the output rows and publishes a row's
    // record once both its source rows are armed.  Consumers report the ROW they are on; the ring sequence number that
    // releases is looked up here (q0s), so sequence arithmetic lives in this warp only.
    if (tid >= ncons) {                      // (the scanner warp, tid >= ncons + 32, left above)
        const uint32_t lane = tid - ncons;
        auto frame_ptr = [&](uint32_t f) -> const uint8_t* {
            return p.in_ptrs ? p.in_ptrs[f] : p.in_base + (unsigned long long)f * p.frame_pitch;
        };
        const size_t plane = (size_t)p.dst_w * p.dst_h;
        // Work ranges [first, end) of global output rows, in the order this CTA takes them: its static run, then chunks
        // from the launch's counter.  The arming cursor fetches them; the record cursor, which lags it, reads them back
        // from a queue of 8 (more than the ring can hold rows of).  An ITEM is the part of a range inside one frame.
        unsigned long long* rq = reinterpret_cast<unsigned long long*>(smem + kSmemRq);
        uint32_t nfetched = 0;                                   // ranges fetched so far
        bool exhausted = false;                                  // the counter has run out (or there is no dynamic part)
        auto fetch_range = [&](uint32_t ranges_done_by_records) -> bool {
            if (exhausted || nfetched - ranges_done_by_records >= 8) return false;
            unsigned long long ra_ = g0, rb_ = g1;
            if (nfetched > 0) {
                if (p.dyn_chunks == 0) { exhausted = true; return false; }
                uint32_t c = 0;
                if (lane == 0) c = atomicAdd(&g_pre_work[p.dyn_slot], 1u);
                c = __shfl_sync(0xffffffffu, c, 0);
                if (c >= p.dyn_chunks) { exhausted = true; return false; }
                ra_ = p.static_rows + (unsigned long long)c * p.dyn_rows;
                rb_ = min(ra_ + p.dyn_rows, total);
            }
            if (lane == 0) { rq[2 * (nfetched & 7)] = ra_; rq[2 * (nfetched & 7) + 1] = rb_; }
            __syncwarp();
            ++nfetched;
            return true;
        };
        // cursor over the ranges: (index of its range, next global row, end of the range)
        uint32_t ia = 0, ir = 0;                                 // ranges the arming / record cursor has finished
        unsigned long long ga = 0, gea = 0, gr = 0, ger = 0;     // current range of each cursor: next row, end
        bool a_has = false, r_has = false;                       // cursor holds a range
        uint32_t fa = 0, seq = 0, a_a = 0;                       // arming cursor: frame, sequences armed so far, first inside row
        int ra = 0, na = 0, ylo_a = 0;                           //   rows armed / to arm in the item, its first source row
        const uint8_t* base_a = nullptr;
        bool a_open = false, a_done = false;
        uint32_t fr = 0, hr = 0, h1r = 0, a_r = 0, b_r = 0, seqbase_r = 0, n_pub = 0;   // record cursor
        int ylo_r = 0, nrows_r = 0;
        bool r_open = false, r_done = false;
#ifdef TK_B200_CHECKED
        volatile uint32_t* chk = reinterpret_cast<volatile uint32_t*>(smem + kSmemChk);
#endif
        for (;;) {
            // ---- open the next item of each cursor ----
            if (!a_open && !a_done) {
                if (a_has && ga == gea) { a_has = false; ++ia; }
                if (!a_has) {
                    if (ia < nfetched || fetch_range(ir)) { ga = rq[2 * (ia & 7)]; gea = rq[2 * (ia & 7) + 1]; a_has = true; }
                    else if (exhausted) a_done = true;
                }
                if (a_has && ga < gea) {
                    fa = (uint32_t)(ga / p.dst_h);
                    const uint32_t h0 = (uint32_t)(ga - (unsigned long long)fa * p.dst_h);
                    const uint32_t h1 = (uint32_t)min((unsigned long long)p.dst_h, h0 + (gea - ga));
                    item_src_rows(p, h0, h1, ylo_a, na, a_a);
                    base_a = frame_ptr(fa); ra = 0; a_open = true;
                    ga += h1 - h0;
                }
            }
            if (!r_open && !r_done) {
                if (r_has && gr == ger) { r_has = false; ++ir; }
                if (!r_has) {
                    if (ir < nfetched) { gr = rq[2 * (ir & 7)]; ger = rq[2 * (ir & 7) + 1]; r_has = true; }
                    else if (a_done) r_done = true;              // the arming cursor fetches first: nothing more will come
                }
                if (r_has && gr < ger) {
                    fr = (uint32_t)(gr / p.dst_h);
                    const uint32_t h0 = (uint32_t)(gr - (unsigned long long)fr * p.dst_h);
                    const uint32_t h1 = (uint32_t)min((unsigned long long)p.dst_h, h0 + (ger - gr));
                    item_src_rows(p, h0, h1, ylo_r, nrows_r, a_r);
                    hr = h0; h1r = h1; b_r = min(h1, p.pad_top + p.new_h); r_open = true;
                    gr += h1 - h0;
                }
            }
            if (r_done) {
                // every row has its record: one END record per row group, then this warp is done
                const uint32_t min_n_ = __reduce_min_sync(0xffffffffu, progress[lane]);
                if ((unsigned long long)n_pub + p.ty <= (unsigned long long)min_n_ + kRecs) {
                    if (lane < p.ty) {
                        const uint32_t e = (n_pub + lane) & (kRecs - 1);
                        recs[2 * e] = make_uint4(p.ring_off, p.ring_off, 0u, kRecEnd);
                        recs[2 * e + 1] = make_uint4(0u, 0u, 0u, 0u);
                        q0s[e] = seqbase_r;
#ifdef TK_B200_CHECKED
                        chk[3 * e] = 0; chk[3 * e + 1] = 0; chk[3 * e + 2] = n_pub + lane;
#endif
                    }
                    __syncwarp();
                    if (lane == 0) { __threadfence_block(); *rec_ready = n_pub + p.ty; }
                    break;
                }
                __nanosleep(64);
                continue;
            }
            // lane j looks at output row hr + j of the record cursor's item
            const uint32_t h = hr + lane;
            const bool in_item = r_open && h < h1r;
            const bool inside = in_item && h >= a_r && h < b_r;
            int y = 0; float yd = 0.0f;
            // padding: above the region nothing is released yet, below it everything; no item open: everything armed so far
            // belongs to rows that are published (seqbase_r already counts the closed item)
            uint32_t q0 = seqbase_r + ((r_open && h >= b_r) ? (uint32_t)nrows_r : 0u);
            if (inside) {
                const float gy = __fmul_rn(p.y_ratio, (float)(h - p.pad_top));    // src/vision/tk_image_preprocessor.c:49,51,53
                y = min(__float2int_rz(gy), (int)p.src_h - 2);
                yd = __fsub_rn(gy, (float)y);
                q0 = seqbase_r + (p.sparse ? 2u * (h - a_r) : (uint32_t)(y - ylo_r));
            }
            // every sequence below min_rel has been released by all consumer warps
            const uint32_t min_n = __reduce_min_sync(0xffffffffu, progress[lane]);
            const uint32_t q0_next = __shfl_sync(0xffffffffu, q0, 0);
            const uint32_t min_rel = min_n < n_pub ? q0s[min_n & (kRecs - 1)] : q0_next;
            bool did = false;
            if (a_open) {
                const uint32_t room = min_rel + S - seq;          // sequences seq .. seq + room - 1 may be armed now
                if constexpr (TMA) {
                    // Rows are armed in BATCHES, one lane per row.  (One row per loop trip — ~95 dependent instructions each,
                    // two CTA fences among them — made this single warp the kernel's pacemaker: ncu showed it busy throughout,
                    // the slot it wanted already free 78 % of the time, 20 % of the consumers' samples on the `full` barriers.)
                    const uint32_t k = min(min(room, 32u), (uint32_t)(na - ra));
                    if (k) {
                        __threadfence_block();                    // the releases read above happen-before the overwrites below
                        if (lane < k) {
                            const int r = ra + (int)lane;
                            const uint32_t q = seq + lane, slot = q & smask;
                            const int row = p.sparse ? src_row_of(p, a_a + ((uint32_t)r >> 1)) + (r & 1) : ylo_a + r;
                            // every staged row is waited on by the warps that sample it before they move past it, so the
                            // slot's previous copy has landed; the check is free then.
                            if (q >= S) mbar_wait(&full[slot], ((q >> p.s_log2) - 1u) & 1u);
                            uint8_t* dst = ring + (size_t)slot * p.slot_bytes;
                            const uint8_t* src = base_a + (size_t)row * p.src_stride;
                            fence_proxy_async();
                            if constexpr (FMT == FMT_I420) {
                                const uint32_t cr = min((uint32_t)row >> 1, p.chroma_h - 1);
                                mbar_arrive_expect_tx(&full[slot], p.copy_bytes + 2 * p.chroma_copy_bytes);
                                bulk_g2s(dst, src, p.copy_bytes, &full[slot]);
                                bulk_g2s(dst + p.u_slot_off, base_a + p.u_off + (size_t)cr * p.chroma_pitch, p.chroma_copy_bytes, &full[slot]);
                                bulk_g2s(dst + p.v_slot_off, base_a + p.v_off + (size_t)cr * p.chroma_pitch, p.chroma_copy_bytes, &full[slot]);
                            } else {
                                mbar_arrive_expect_tx(&full[slot], p.copy_bytes);
                                bulk_g2s(dst, src, p.copy_bytes, &full[slot]);
                            }
                        }
                        __syncwarp();
                        seq += k; ra += (int)k; did = true;
                    }
                } else {
                    if (room && ra < na) {
                        __threadfence_block();
                        const int row = p.sparse ? src_row_of(p, a_a + ((uint32_t)ra >> 1)) + (ra & 1) : ylo_a + ra;
                        const uint32_t slot = seq & smask;
                        uint8_t* dst = ring + (size_t)slot * p.slot_bytes;
                        const uint8_t* src = base_a + (size_t)row * p.src_stride;
                        const uint32_t nb = (FMT == FMT_RGB8) ? p.src_w * 3 : (FMT == FMT_RGBA8 ? p.src_w * 4 : p.src_w);
                        warp_copy_unaligned(dst, src, nb, lane);
                        if constexpr (FMT == FMT_I420) {
                            const uint32_t cr = min((uint32_t)row >> 1, p.chroma_h - 1);
                            warp_copy_unaligned(dst + p.u_slot_off, base_a + p.u_off + (size_t)cr * p.chroma_pitch, p.src_w >> 1, lane);
                            warp_copy_unaligned(dst + p.v_slot_off, base_a + p.v_off + (size_t)cr * p.chroma_pitch, p.src_w >> 1, lane);
                        }
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&full[slot]);
                        __syncwarp();
                        seq += 1; ra += 1; did = true;
                    }
                }
                if (ra == na) a_open = false;
            }
            // Publish the records of the leading rows whose two source rows are armed by now (a waiter tells barrier
            // phases apart by parity only, so it may look at full[slot] for sequence q only once q has been armed, which
            // implies q - S has completed) and whose entry is free: row n - kRecs is behind every consumer warp.
            {
                const bool ok = in_item && (unsigned long long)n_pub + lane < (unsigned long long)min_n + kRecs &&
                                (!inside || q0 + 1 < seq);
                const uint32_t bal = __ballot_sync(0xffffffffu, ok);
                const uint32_t cnt = bal == 0xffffffffu ? 32u : (uint32_t)__ffs((int)~bal) - 1u;
                if (lane < cnt) {
                    const uint32_t e = (n_pub + lane) & (kRecs - 1);
                    const uint32_t s0 = q0 & smask, s1 = (q0 + 1) & smask;
                    uint4 r0, r1;
                    r0.x = p.ring_off + s0 * p.slot_bytes; r0.y = p.ring_off + s1 * p.slot_bytes;
                    r0.z = __float_as_uint(yd);
                    r0.w = (inside ? 1u : 0u) | (((q0 >> p.s_log2) & 1u) << 1) | ((((q0 + 1) >> p.s_log2) & 1u) << 2);
                    r1.x = smem_u32(&full[s0]); r1.y = smem_u32(&full[s1]);
                    const unsigned long long op = reinterpret_cast<unsigned long long>(p.out) +
                                                  ((unsigned long long)fr * 3 * plane + (unsigned long long)h * p.dst_w) * ESZ;
                    r1.z = (uint32_t)op; r1.w = (uint32_t)(op >> 32);
                    recs[2 * e] = r0; recs[2 * e + 1] = r1;
                    q0s[e] = q0;
#ifdef TK_B200_CHECKED
                    chk[3 * e] = fr; chk[3 * e + 1] = (uint32_t)y; chk[3 * e + 2] = n_pub + lane;
#endif
                }
                __syncwarp();
                if (cnt) {
                    if (lane == 0) { __threadfence_block(); *rec_ready = n_pub + cnt; }
                    n_pub += cnt; hr += cnt; did = true;
                    if (hr == h1r) { r_open = false; seqbase_r += (uint32_t)nrows_r; }
                }
            }
            if (!did) __nanosleep(64);
        }
        if (p.dyn_chunks && lane == 0) {
            // the last CTA of the launch to get here leaves the slot's counters at zero for the launch that takes it next
            __threadfence();
            if (atomicAdd(&g_pre_done[p.dyn_slot], 1u) == gridDim.x - 1) {
                g_pre_work[p.dyn_slot] = 0u; g_pre_done[p.dyn_slot] = 0u;
                __threadfence();
            }
        }
        finish();
        return;
    }

    // ======================= consumer warps ======================================
    // Warp (cy, *) produces the CTA's rows cy, cy + ty, ...  Before a row it publishes the row number (everything its
    // earlier rows read is released), takes the row's record, waits for the two source rows on their `full` barriers
    // and computes.  There is no CTA-wide barrier after set-up.
    const uint32_t cx = tid % p.tx, cy = tid / p.tx, lane = tid & 31;

    // Pad rows/pixels (letterbox) carry the same normalisation as real pixels.
    float padn[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) padn[c] = norm_only<EXACT>(p, c, p.pad_value);

    const uint32_t Q = p.tab_w >> 2;
#ifdef TK_B200_CHECKED
    volatile uint32_t* chk = reinterpret_cast<volatile uint32_t*>(smem + kSmemChk);
    auto frame_ptr = [&](uint32_t f) -> const uint8_t* {
        return p.in_ptrs ? p.in_ptrs[f] : p.in_base + (unsigned long long)f * p.frame_pitch;
    };
#endif

    // Column constants of a quad of four output pixels (row-independent).
    struct Cols {
        int off[4]; uint32_t aux[4];             // col_consts
        uint32_t padmask;                        // letterbox: bit j = pixel j is padding
        float2 xd2[2], omx2[2], cxd2[2], comx2[2];   // fp32 arithmetic: xd, 1 - xd, and (EXACT) their -2^23 multiples, as packed pairs
        uint32_t xdh[4];                         // HALF: half2 {xd, xd}
    };
    auto load_cols = [&](uint32_t q, Cols& c) {
        const int4 xb4 = reinterpret_cast<const int4*>(xb_tab)[q];
        const int xbs[4] = {xb4.x, xb4.y, xb4.z, xb4.w};
        c.padmask = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (LBOX && xbs[j] < 0) c.padmask |= 1u << j;
            col_consts<FMT>(p, LBOX ? max(xbs[j], 0) : xbs[j], c.off[j], c.aux[j]);
        }
        if constexpr (AR == AR_HALF) {
            const uint4 xd4 = reinterpret_cast<const uint4*>(xd_tab)[q];
            c.xdh[0] = xd4.x; c.xdh[1] = xd4.y; c.xdh[2] = xd4.z; c.xdh[3] = xd4.w;
        } else {
            const float4 xd4 = reinterpret_cast<const float4*>(xd_tab)[q];
            c.xd2[0] = f2(xd4.x, xd4.y); c.xd2[1] = f2(xd4.z, xd4.w);
#pragma unroll
            for (int jp = 0; jp < 2; ++jp) {
                c.omx2[jp] = add2(f2s(1.0f), f2(-c.xd2[jp].x, -c.xd2[jp].y));      // 1 - xd
                c.cxd2[jp] = c.xd2[jp]; c.comx2[jp] = c.omx2[jp];
                if constexpr (EXACT) { c.cxd2[jp] = mul2(c.xd2[jp], f2s(-kBias)); c.comx2[jp] = mul2(c.omx2[jp], f2s(-kBias)); }
            }
        }
    };

    uint32_t ready_seen = 0;                     // records known to be published (acquire-loaded from rec_ready)
    // Row set-up: publish the row number, take the record, wait for the two source rows.
    struct Row { const uint8_t* row0; const uint8_t* row1; uint8_t* out; float yd, omy; bool vin, end;
#ifdef TK_B200_CHECKED
                 uint32_t chk_f; int chk_y;
#endif
    };
    auto begin_row = [&](uint32_t n, Row& r) {
        __syncwarp();                            // all lanes are done reading the rows being released
        if (lane == 0) sts_release(&progress[tid >> 5], n);
        // the producer runs up to kRecs records ahead: most rows find their record published by an earlier look
        if (n >= ready_seen) {
            while ((ready_seen = lds_acquire(rec_ready)) <= n) __nanosleep(32);
        }
        const uint32_t e = n & (kRecs - 1);
        const uint4 ra = recs[2 * e], rb = recs[2 * e + 1];
        r.end = (ra.w & kRecEnd) != 0;           // no more rows for this CTA (the producer's last records)
        r.vin = !r.end && (!LBOX || (ra.w & 1u));
        r.yd = __uint_as_float(ra.z); r.omy = __fsub_rn(1.0f, r.yd);
        r.row0 = smem + ra.x; r.row1 = smem + ra.y;
        r.out = reinterpret_cast<uint8_t*>(((unsigned long long)rb.w << 32) | rb.z);       // output row, tensor plane 0
        if (r.vin) {
            mbar_wait_backoff_addr(rb.x, (ra.w >> 1) & 1u);
            mbar_wait_backoff_addr(rb.y, (ra.w >> 2) & 1u);
        }
#ifdef TK_B200_CHECKED
        TK_DEV_ASSERT(chk[3 * e + 2] == n && n < *rec_ready);                            // the record is this row's
        TK_DEV_ASSERT(ra.x >= p.ring_off && ra.x + p.slot_bytes <= p.ring_off + S * p.slot_bytes && (ra.x - p.ring_off) % p.slot_bytes == 0);
        TK_DEV_ASSERT(ra.y >= p.ring_off && ra.y + p.slot_bytes <= p.ring_off + S * p.slot_bytes && (ra.y - p.ring_off) % p.slot_bytes == 0);
        r.chk_f = chk[3 * e]; r.chk_y = (int)chk[3 * e + 1];
        if (r.vin) {
            check_ring_row<FMT>(p, frame_ptr(r.chk_f), r.chk_y, r.row0, lane);
            check_ring_row<FMT>(p, frame_ptr(r.chk_f), r.chk_y + 1, r.row1, lane);
        }
#endif
    };
    auto end_row = [&](const Row& r) {
#ifdef TK_B200_CHECKED
        // ... and still hold them now that every pixel of the row has been computed (nothing overwrote a slot early)
        if (r.vin) {
            check_ring_row<FMT>(p, frame_ptr(r.chk_f), r.chk_y, r.row0, lane);
            check_ring_row<FMT>(p, frame_ptr(r.chk_f), r.chk_y + 1, r.row1, lane);
        }
#endif
    };

    // Four consecutive output pixels x three channels of one row.
    auto do_quad = [&](uint32_t q, const Cols& c, const Row& r) {
        if constexpr (AR == AR_HALF) {
            const uint32_t pad_rg = h2_bits(__hfma2(__float2half2_rn(p.pad_value - 128.0f), as_h2(p.h_scale_rg), as_h2(p.h_bias_rg)));
            const uint32_t pad_bb = h2_bits(__hfma2(__float2half2_rn(p.pad_value - 128.0f), as_h2(p.h_scale_bb), as_h2(p.h_bias_bb)));
            const __half2 yd2 = __float2half2_rn(r.yd);
            uint32_t org[4], obb[4];
            if (r.vin) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (LBOX && ((c.padmask >> j) & 1u)) { org[j] = pad_rg; obb[j] = pad_bb; continue; }
                    uint32_t tl_rg, tl_bb, tr_rg, tr_bb, bl_rg, bl_bb, br_rg, br_bb;
                    fetch_pairs_half<FMT>(r.row0, c.off[j], c.aux[j], tl_rg, tl_bb, tr_rg, tr_bb);
                    fetch_pairs_half<FMT>(r.row1, c.off[j], c.aux[j], bl_rg, bl_bb, br_rg, br_bb);
                    const __half2 xd2 = as_h2(c.xdh[j]);
                    org[j] = h2_bits(__hfma2(lerp_pair_half(tl_rg, tr_rg, bl_rg, br_rg, xd2, yd2), as_h2(p.h_scale_rg), as_h2(p.h_bias_rg)));
                    obb[j] = h2_bits(__hfma2(lerp_pair_half(tl_bb, tr_bb, bl_bb, br_bb, xd2, yd2), as_h2(p.h_scale_bb), as_h2(p.h_bias_bb)));
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) { org[j] = pad_rg; obb[j] = pad_bb; }
            }
            // regroup {R,G}/{B,B} per pixel into per-plane pixel pairs
            const uint2 vr = make_uint2(__byte_perm(org[0], org[1], 0x5410), __byte_perm(org[2], org[3], 0x5410));
            const uint2 vg = make_uint2(__byte_perm(org[0], org[1], 0x7632), __byte_perm(org[2], org[3], 0x7632));
            const uint2 vb = make_uint2(__byte_perm(obb[0], obb[1], 0x5410), __byte_perm(obb[2], obb[3], 0x5410));
            uint8_t* qp = r.out + (size_t)(4u * q) * 2;
            const uint2 vs[3] = {vr, vg, vb};
            if (p.vec_store) {
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) __stcs(reinterpret_cast<uint2*>(qp + p.cplane[ch]), vs[ch]);
            } else {
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    const uint32_t wds[2] = {vs[ch].x, vs[ch].y};
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (4u * q + j < p.dst_w)
                            reinterpret_cast<uint16_t*>(qp + p.cplane[ch])[j] = (uint16_t)(wds[j >> 1] >> ((j & 1) * 16));
                }
            }
        } else {
            float o[3][4];
            if (r.vin) {
#pragma unroll
                for (int j = 0; j < 4; j += 2) {                 // two neighbouring pixels per packed-fp32 pass
                    float ma[4][3], mb[4][3];
                    constexpr bool XU = kPreXu && EXACT && FMT == FMT_RGB8;
                    fetch_quad<FMT, XU>(p, r.row0, r.row1, c.off[j], c.aux[j], ma);
                    fetch_quad<FMT, XU>(p, r.row0, r.row1, c.off[j + 1], c.aux[j + 1], mb);
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        const float2 v = blend_norm2<EXACT>(p, ch, f2(ma[0][ch], mb[0][ch]), f2(ma[1][ch], mb[1][ch]),
                                                            f2(ma[2][ch], mb[2][ch]), f2(ma[3][ch], mb[3][ch]), c.xd2[j >> 1],
                                                            c.omx2[j >> 1], c.cxd2[j >> 1], c.comx2[j >> 1], r.yd, r.omy,
                                                            XU && ch == 2, XU);
                        o[ch][j] = (LBOX && ((c.padmask >> j) & 1u)) ? padn[ch] : v.x;
                        o[ch][j + 1] = (LBOX && ((c.padmask >> (j + 1)) & 1u)) ? padn[ch] : v.y;
                    }
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) { o[0][j] = padn[0]; o[1][j] = padn[1]; o[2][j] = padn[2]; }
            }
            // one 64-bit address per quad; the channel planes are constant byte offsets from it
            uint8_t* qp = r.out + (size_t)(4u * q) * ESZ;
            if (p.vec_store) {
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
                    if constexpr (F16) {
                        __half2 lo = __floats2half2_rn(o[ch][0], o[ch][1]);
                        __half2 hi = __floats2half2_rn(o[ch][2], o[ch][3]);
                        uint2 v;
                        v.x = *reinterpret_cast<uint32_t*>(&lo);
                        v.y = *reinterpret_cast<uint32_t*>(&hi);
                        __stcs(reinterpret_cast<uint2*>(qp + p.cplane[ch]), v);
                    } else {
                        __stcs(reinterpret_cast<float4*>(qp + p.cplane[ch]), make_float4(o[ch][0], o[ch][1], o[ch][2], o[ch][3]));
                    }
                }
            } else {
#pragma unroll
                for (int ch = 0; ch < 3; ++ch) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        if (4u * q + j < p.dst_w) {
                            if constexpr (F16) reinterpret_cast<__half*>(qp + p.cplane[ch])[j] = __float2half_rn(o[ch][j]);
                            else reinterpret_cast<float*>(qp + p.cplane[ch])[j] = o[ch][j];
                        }
                    }
                }
            }
        }
    };

    uint32_t rows_done = 0;
    if (Q <= p.tx) {
        // One quad per thread and row (rows of <= 640 pixels): its column constants stay in registers for the whole run —
        // two 16-byte table loads, six packed fp32 and a dozen ALU-pipe instructions less per row.
        Cols c{};
        const bool has = cx < Q;
        if (has) load_cols(cx, c);
        for (uint32_t n = cy;; n += p.ty) {
            Row r; begin_row(n, r);
            if (r.end) break;
            if (has) do_quad(cx, c, r);
            end_row(r);
            ++rows_done;
        }
    } else {
        for (uint32_t n = cy;; n += p.ty) {
            Row r; begin_row(n, r);
            if (r.end) break;
            ++rows_done;
            for (uint32_t q = cx; q < Q; q += p.tx) {
                Cols c{};
                load_cols(q, c);
                do_quad(q, c, r);
            }
            end_row(r);
        }
    }
    __syncwarp();                                // done: nothing this warp could still read
    if (lane == 0) { __threadfence_block(); progress[tid >> 5] = 0xffffffffu; }
    if (p.trace.buf && tid == 0) { t_trace[1] = gtimer(); trace_put(p.trace, 1u, blockIdx.x, rows_done, t_trace, 2); }
    finish();
}

// ------------------------------------------------------------------------- //
// Host side                                                                   //
// ------------------------------------------------------------------------- //

// Correctly rounded fp32 reciprocal (needed by the Markstein divide).
static float rn_reciprocal(float s) {
    float y = (float)(1.0 / (double)s);
    float best = y;
    double best_err = 1e300;
    for (int d = -1; d <= 1; ++d) {
        uint32_t b = tk_dm_float_to_bits(y) + (uint32_t)d;
        float c = tk_dm_bits_to_float(b);
        double err = 1.0 - (double)s * (double)c;      // exact product, exact near 1
        if (err < 0) err = -err;
        if (err < best_err) { best_err = err; best = c; }
    }
    return best;
}

void letterbox_geometry(uint32_t sw, uint32_t sh, uint32_t dw, uint32_t dh,
                        uint32_t* nw, uint32_t* nh, uint32_t* pl, uint32_t* pt) {
    unsigned long long a = (unsigned long long)dw * sh, b = (unsigned long long)dh * sw;
    if (a <= b) {
        *nw = dw;
        unsigned long long v = (2ull * sh * dw + sw) / (2ull * sw);
        *nh = (uint32_t)(v < 1 ? 1 : (v > dh ? dh : v));
    } else {
        *nh = dh;
        unsigned long long v = (2ull * sw * dh + sh) / (2ull * sh);
        *nw = (uint32_t)(v < 1 ? 1 : (v > dw ? dw : v));
    }
    *pl = (dw - *nw) / 2;
    *pt = (dh - *nh) / 2;
}

template <int FMT, bool F16, int AR, bool LBOX, bool TMA>
static tk_error_code_t launch_variant(const PreK& k, dim3 grid, dim3 block, size_t smem, cudaStream_t stream) {
    auto kern = preprocess_kernel<FMT, F16, AR, LBOX, TMA>;
    TK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), TK_ERROR_GPU_CUDA_ERROR);
    // Largest shared-memory carve-out, like post_kernel (post.cu), which the pipeline runs beside this kernel: with one
    // L1 / shared split for both, an SM never drains to reconfigure between them.
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    // k.pdl (set only by tk_b200_pipeline_process_device, never by the public launchers): launched as a programmatic
    // dependent.  The kernel reads the caller's frames BEFORE its griddepcontrol.wait, so this is only sound behind a
    // kernel that does not write them and has announced (griddepcontrol.launch_dependents) that its dependents need not
    // wait for it — nms_kernel of the pipeline's previous step, which keeps 64 of 148 SMs busy at the bench batch: the
    // pre-process CTAs take the idle SMs at once (150.9 -> 147.6 us per 64-frame step, bit-exact).  Every thread ends on
    // griddepcontrol.wait, so the grid completes after that kernel and stream order still holds for whatever follows.
    // Behind copies, events or kernels without the trigger (whatever the caller enqueued in between) nothing changes:
    // those complete, memory and all, before this grid starts.  TK_B200_PDL_PRE=0 turns it off.
    static const bool pdl_env = [] { const char* e = getenv("TK_B200_PDL_PRE"); return !e || atoi(e) != 0; }();
    if (k.pdl && pdl_env) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        TK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, k), TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE);
    } else {
        kern<<<grid, block, smem, stream>>>(k);
    }
    return launch_status("preprocess_kernel");
}

template <int FMT, bool F16, int AR>
static tk_error_code_t launch_fmt(const PreK& k, bool lbox, bool tma, dim3 g, dim3 b, size_t s, cudaStream_t st) {
    if (lbox) return tma ? launch_variant<FMT, F16, AR, true, true>(k, g, b, s, st)
                         : launch_variant<FMT, F16, AR, true, false>(k, g, b, s, st);
    return tma ? launch_variant<FMT, F16, AR, false, true>(k, g, b, s, st)
               : launch_variant<FMT, F16, AR, false, false>(k, g, b, s, st);
}

template <int FMT>
static tk_error_code_t launch_type(const PreK& k, bool f16, int ar, bool lbox, bool tma, dim3 g, dim3 b, size_t s, cudaStream_t st) {
    if constexpr (FMT != FMT_I420) {
        if (f16 && ar == AR_HALF) return launch_fmt<FMT, true, AR_HALF>(k, lbox, tma, g, b, s, st);
    }
    if (ar == AR_HALF) ar = AR_FAST;                   // I420 or fp32 output: the fp32 lerp
    if (f16) return ar == AR_EXACT ? launch_fmt<FMT, true, AR_EXACT>(k, lbox, tma, g, b, s, st)
                                   : launch_fmt<FMT, true, AR_FAST>(k, lbox, tma, g, b, s, st);
    return ar == AR_EXACT ? launch_fmt<FMT, false, AR_EXACT>(k, lbox, tma, g, b, s, st)
                          : launch_fmt<FMT, false, AR_FAST>(k, lbox, tma, g, b, s, st);
}

}  // namespace tkb

using namespace tkb;

// True when a launch with these parameters (a) stages only the two source rows of each output row (y_ratio >= 2, so
// a large part of every frame is never fetched) and (b) moves rows with 16-byte bulk copies that end exactly at the
// row's last byte — the conditions under which the pipeline lets the kernel read page-locked HOST frames in place
// over PCIe instead of copying them to HBM first.
bool tkb::preprocess_reads_rows_sparsely(const tk_b200_preprocess_batch_params_t* prm) {
    if (!prm || prm->d_input_ptrs || prm->output_height == 0 || prm->output_width == 0) return false;
    const uint32_t fmt = prm->input_format == TK_PIXEL_FORMAT_RGBA8 ? FMT_RGBA8 : (prm->input_format == TK_PIXEL_FORMAT_I420 ? FMT_I420 : FMT_RGB8);
    const uint32_t bpp = fmt == FMT_RGB8 ? 3 : (fmt == FMT_RGBA8 ? 4 : 1);
    const unsigned long long row_bytes = (unsigned long long)prm->input_width * bpp;
    const unsigned long long stride = prm->input_stride_bytes ? prm->input_stride_bytes : row_bytes;
    if (row_bytes % 16 || stride % 16 || reinterpret_cast<uintptr_t>(prm->d_input) % 16) return false;
    if (fmt == FMT_I420 && ((prm->input_width / 2) % 16 || (stride / 2) % 16 || (stride * prm->input_height) % 16 ||
                            ((stride / 2) * (prm->input_height / 2)) % 16))
        return false;
    if (prm->input_frame_pitch_bytes % 16) return false;
    uint32_t new_w = prm->output_width, new_h = prm->output_height, pl = 0, pt = 0;
    if (prm->resize_mode == TK_B200_RESIZE_LETTERBOX)
        letterbox_geometry(prm->input_width, prm->input_height, prm->output_width, prm->output_height, &new_w, &new_h, &pl, &pt);
    return ((float)prm->input_height - 1.0f) / (float)new_h >= 2.0f;
}

extern "C" tk_error_code_t tk_b200_preprocess_batch(const tk_b200_preprocess_batch_params_t* prm, cudaStream_t stream) {
    return tkb::preprocess_batch_impl(prm, stream, false, nullptr, nullptr);
}

tk_error_code_t tkb::preprocess_batch_impl(const tk_b200_preprocess_batch_params_t* prm, cudaStream_t stream, bool pdl_after_nms,
                                           const tk_b200_decode_params_t* scan, uint32_t* d_scan_bitmap) {
    if (!prm || (!prm->d_input && !prm->d_input_ptrs) || !prm->d_output) return TK_ERROR_INVALID_ARGUMENT;
    if (prm->num_frames == 0) return TK_SUCCESS;
    if (prm->input_width == 0 || prm->input_height == 0 || prm->output_width == 0 || prm->output_height == 0)
        return TK_ERROR_INVALID_DIMENSIONS;
    // The reference mapping reads pixel x+1 / row y+1 (tk_image_preprocessor.c:59-61):
    // undefined for a 1-pixel-wide or 1-row source, rejected here.
    if (prm->input_width < 2 || prm->input_height < 2) return TK_ERROR_INVALID_DIMENSIONS;
    for (int c = 0; c < 3; ++c) {
        float s = (&prm->std_dev.x)[c];
        if (!(s > 0.0f) || !(s < 3.0e38f)) return TK_ERROR_INVALID_ARGUMENT;
    }
    int fmt;
    uint32_t bpp;
    switch (prm->input_format) {
        case TK_PIXEL_FORMAT_RGB8:  fmt = FMT_RGB8;  bpp = 3; break;
        case TK_PIXEL_FORMAT_RGBA8: fmt = FMT_RGBA8; bpp = 4; break;
        case TK_PIXEL_FORMAT_I420:  fmt = FMT_I420;  bpp = 1; break;
        case TK_PIXEL_FORMAT_BGR8:  fmt = FMT_RGB8;  bpp = 3; break;   // same gather, planes swapped on store
        default: return TK_ERROR_INVALID_ARGUMENT;
    }
    const bool bgr = prm->input_format == TK_PIXEL_FORMAT_BGR8;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;

    PreK k{};
    k.in_base = static_cast<const uint8_t*>(prm->d_input);
    k.in_ptrs = reinterpret_cast<const uint8_t* const*>(prm->d_input_ptrs);
    k.out = prm->d_output;
    k.num_frames = prm->num_frames;
    k.src_w = prm->input_width; k.src_h = prm->input_height;
    const uint32_t row_bytes = prm->input_width * bpp;
    k.src_stride = prm->input_stride_bytes ? prm->input_stride_bytes : row_bytes;
    if (k.src_stride < row_bytes) return TK_ERROR_INVALID_ARGUMENT;
    k.dst_w = prm->output_width; k.dst_h = prm->output_height;
    k.new_w = k.dst_w; k.new_h = k.dst_h; k.pad_left = 0; k.pad_top = 0;
    const bool lbox = prm->resize_mode == TK_B200_RESIZE_LETTERBOX;
    if (lbox) letterbox_geometry(k.src_w, k.src_h, k.dst_w, k.dst_h, &k.new_w, &k.new_h, &k.pad_left, &k.pad_top);
    // src/vision/tk_image_preprocessor.c:43-44 (uint -> float, fp32 divide)
    k.x_ratio = ((float)k.src_w - 1.0f) / (float)k.new_w;
    k.y_ratio = ((float)k.src_h - 1.0f) / (float)k.new_h;
    k.swap_rb = bgr ? 1u : 0u;
    k.one = 1.0f;
    for (int c = 0; c < 3; ++c)
        k.cplane[c] = (unsigned long long)(bgr ? 2 - c : c) * k.dst_w * k.dst_h * (prm->output_type == TK_B200_OUT_F16 ? 2u : 4u);
    k.trace = trace_target();
    k.pdl = pdl_after_nms ? 1u : 0u;
    k.trigger = pdl_chain_hint() ? 1u : 0u;
    if (scan && d_scan_bitmap && scan->d_logits && scan->num_proposals) {
        k.scan.logits = scan->d_logits; k.scan.bitmap = d_scan_bitmap;
        k.scan.P = scan->num_proposals; k.scan.L = 5 + scan->num_classes; k.scan.nwf = (scan->num_proposals + 31) / 32;
        const unsigned long long tw = (unsigned long long)scan->num_frames * k.scan.nwf;
        if (tw > 0xffffffffull) return TK_ERROR_INVALID_DIMENSIONS;
        k.scan.total_words = (uint32_t)tw;
        k.scan.thr = scan->confidence_threshold; k.scan.act = (int)scan->activation;
    }
    for (int c = 0; c < 3; ++c) {
        // constants are indexed by SOURCE channel; BGR byte c is tensor channel 2 - c
        const int tc = bgr ? 2 - c : c;
        float mu = (&prm->mean.x)[tc], sd = (&prm->std_dev.x)[tc];
        k.mean[c] = mu; k.stdv[c] = sd; k.rstd[c] = rn_reciprocal(sd);
        k.fscale[c] = (float)(1.0 / (255.0 * (double)sd));
        k.fbias[c] = (float)(-(double)mu / (double)sd);
    }
    k.pad_value = prm->pad_value;
    {
        auto h2 = [](float lo, float hi) {
            __half2 v = __floats2half2_rn(lo, hi);
            uint32_t u; memcpy(&u, &v, 4); return u;
        };
        // HALF works on byte - 128: bias' = -mean/std + 128/(255 std), rounded once from double
        float hb[3];
        for (int c = 0; c < 3; ++c) {
            const int tc = bgr ? 2 - c : c;
            const double mu = (&prm->mean.x)[tc], sd = (&prm->std_dev.x)[tc];
            hb[c] = (float)(-mu / sd + 128.0 / (255.0 * sd));
        }
        k.h_scale_rg = h2(k.fscale[0], k.fscale[1]); k.h_bias_rg = h2(hb[0], hb[1]);
        k.h_scale_bb = h2(k.fscale[2], k.fscale[2]); k.h_bias_bb = h2(hb[2], hb[2]);
    }
    unsigned long long frame_bytes = (unsigned long long)k.src_stride * k.src_h;
    if (fmt == FMT_I420) {
        if (k.src_w < 4 || k.src_h < 4) return TK_ERROR_INVALID_DIMENSIONS;
        k.chroma_pitch = k.src_stride / 2;
        k.chroma_h = k.src_h / 2;
        k.u_off = (unsigned long long)k.src_stride * k.src_h;
        k.v_off = k.u_off + (unsigned long long)k.chroma_pitch * k.chroma_h;
        frame_bytes = k.v_off + (unsigned long long)k.chroma_pitch * k.chroma_h;
    }
    k.frame_pitch = prm->input_frame_pitch_bytes ? prm->input_frame_pitch_bytes : frame_bytes;

    // Ring slot layout and the TMA eligibility test (16-byte rule of cp.async.bulk).
    const uint32_t luma_slot = (row_bytes + 8 + 15) & ~15u;      // +8: the 3-word gather may overrun by <= 6 B
    uint32_t slot = luma_slot;
    if (fmt == FMT_I420) {
        const uint32_t cbytes = ((k.src_w / 2) + 15) & ~15u;
        k.u_slot_off = luma_slot; k.v_slot_off = luma_slot + cbytes;
        slot = luma_slot + 2 * cbytes;
        k.chroma_copy_bytes = cbytes;
    }
    k.slot_bytes = slot;
    k.copy_bytes = (row_bytes + 15) & ~15u;
    bool tma = !prm->d_input_ptrs && (reinterpret_cast<uintptr_t>(prm->d_input) % 16 == 0) &&
               (k.frame_pitch % 16 == 0) && (k.src_stride % 16 == 0) && (k.copy_bytes <= k.src_stride);
    if (fmt == FMT_I420)
        tma = tma && (k.u_off % 16 == 0) && (k.v_off % 16 == 0) && (k.chroma_pitch % 16 == 0) &&
              (k.chroma_copy_bytes <= k.chroma_pitch);

    k.tab_w = (k.dst_w + 3) & ~3u;
    k.ring_off = (kSmemTabs + 8 * k.tab_w + 127) & ~127u;
    k.vec_store = (k.dst_w % 4 == 0) && (reinterpret_cast<uintptr_t>(prm->d_output) % 16 == 0);

    // Consumer shape: tx threads along x cover the row's pixel quads, ty rows in flight;
    // one more warp produces.
    const uint32_t Q = k.tab_w / 4;
    uint32_t tx = Q >= 160 ? 160 : ((Q + 31) & ~31u);
    uint32_t ty = 320 / tx; if (ty < 1) ty = 1;

    // Ring depth: the largest power of two (<= 32 slots) that fits the budget — two CTAs
    // per SM when that still covers the source rows `ty` consecutive output rows span plus
    // two rows of look-ahead; otherwise fewer rows in flight, then one CTA per SM.
    k.sparse = k.y_ratio >= 2.0f ? 1u : 0u;
    auto span_rows = [&](uint32_t g) { return k.sparse ? 2u * g : (uint32_t)((double)(g - 1) * (double)k.y_ratio) + 3; };
    uint32_t rows_in_flight = 0, s_log2 = 0;
    const size_t budgets[2] = {100 * 1024, (size_t)dev->max_smem_optin};
    for (int bi = 0; bi < 2 && !rows_in_flight; ++bi) {
        const size_t avail = budgets[bi] > k.ring_off ? budgets[bi] - k.ring_off : 0;
        const uint32_t slots = (uint32_t)(avail / slot);
        if (slots < 2) continue;
        uint32_t lg = 0;
        while ((2u << lg) <= slots && lg < 5) ++lg;
        for (uint32_t g = ty; g >= 1; --g) {
            const uint32_t gg = g < k.dst_h ? g : k.dst_h;
            if ((1u << lg) >= span_rows(gg) + (bi == 0 && g > 1 ? 2u : 0u)) { rows_in_flight = gg; s_log2 = lg; break; }
        }
    }
    if (!rows_in_flight) {
        tk_error_set_detail("source rows of %u bytes do not fit the shared-memory ring", slot);
        return TK_ERROR_NOT_IMPLEMENTED;
    }
    k.s_log2 = s_log2;
    k.tx = tx; k.ty = rows_in_flight;
    const size_t smem = k.ring_off + ((size_t)slot << s_log2);
    dim3 block(tx * rows_in_flight + 32 + (k.scan.logits ? 32 : 0));      // consumers + producer warp (+ scanner warp)

    const unsigned long long total_rows = (unsigned long long)k.num_frames * k.dst_h;
    const unsigned long long per_cta_min = 2ull * rows_in_flight;            // at least two rows per consumer stream
    unsigned long long max_ctas = (total_rows + per_cta_min - 1) / per_cta_min;
    const uint32_t resident = (smem <= 110 * 1024 ? 2u : 1u) * (uint32_t)dev->sm_count;
    dim3 grid((unsigned)(max_ctas < resident ? max_ctas : resident));
    if (total_rows / grid.x >= 0x7fffffffull) {                               // the kernel counts a CTA's rows in 32 bits
        tk_error_set_detail("%llu output rows are more than one launch takes", total_rows);
        return TK_ERROR_NOT_IMPLEMENTED;
    }

    // Static / dynamic split (PreK): the dynamic part only when a CTA has enough rows for it to matter.
    // TK_B200_PRE_DYN="<percent handed out dynamically>,<rows per chunk>" (experiments; "0" = all static).
    struct DynCfg { uint32_t pct, rows; };
    static const DynCfg dyn = [] {
        DynCfg c{30, 4};
        if (const char* e = getenv("TK_B200_PRE_DYN")) {
            unsigned a = 0, b = 0;
            const int got = sscanf(e, "%u,%u", &a, &b);
            if (got >= 1) c.pct = a > 100 ? 100 : a;
            if (got >= 2 && b >= 1 && b <= 64) c.rows = b;
        }
        return c;
    }();
    k.static_rows = total_rows; k.dyn_rows = 0; k.dyn_chunks = 0; k.dyn_slot = 0;
    if (dyn.pct && total_rows / grid.x >= 32) {
        const uint32_t cr = dyn.rows * rows_in_flight;                       // chunk: dyn.rows rows per consumer row group
        unsigned long long dyn_total = total_rows * dyn.pct / 100;
        unsigned long long chunks = dyn_total / cr;
        if (chunks > 0x7fffffffull) chunks = 0x7fffffffull;
        if (chunks) {
            static std::atomic<uint32_t> next_slot{0};
            k.dyn_rows = cr; k.dyn_chunks = (uint32_t)chunks;
            k.static_rows = total_rows - chunks * cr;
            k.dyn_slot = next_slot.fetch_add(1, std::memory_order_relaxed) % kPreSlots;
        }
    }

    const bool f16 = prm->output_type == TK_B200_OUT_F16;
    if ((unsigned)prm->arith > TK_B200_ARITH_HALF) return TK_ERROR_INVALID_ARGUMENT;
    const int ar = (int)prm->arith;
    switch (fmt) {
        case FMT_RGB8:  return launch_type<FMT_RGB8>(k, f16, ar, lbox, tma, grid, block, smem, stream);
        case FMT_RGBA8: return launch_type<FMT_RGBA8>(k, f16, ar, lbox, tma, grid, block, smem, stream);
        default:        return launch_type<FMT_I420>(k, f16, ar, lbox, tma, grid, block, smem, stream);
    }
}

// src/gpu/cuda/tk_cuda_kernels.cu:252-303 (validation order and codes kept).
extern "C" tk_error_code_t tk_kernels_preprocess_image(const tk_preprocess_params_t* params, cudaStream_t stream) {
    if (!params || !params->d_input_image || !params->d_output_tensor) return TK_ERROR_INVALID_ARGUMENT;
    if (params->input_width == 0 || params->input_height == 0 || params->output_width == 0 || params->output_height == 0)
        return TK_ERROR_INVALID_DIMENSIONS;
    tk_b200_preprocess_batch_params_t b{};
    b.d_input = params->d_input_image;
    b.num_frames = 1;
    b.input_width = params->input_width;
    b.input_height = params->input_height;
    b.input_stride_bytes = params->input_stride_bytes;
    b.input_format = TK_PIXEL_FORMAT_RGB8;
    b.d_output = params->d_output_tensor;
    b.output_width = params->output_width;
    b.output_height = params->output_height;
    b.output_type = TK_B200_OUT_F32;
    b.resize_mode = TK_B200_RESIZE_STRETCH;
    b.arith = TK_B200_ARITH_EXACT;
    b.mean = params->mean;
    b.std_dev = params->std_dev;
    return tk_b200_preprocess_batch(&b, stream);
}
