/*
 * tk_oracle.c — CPU restatement of the reference's vision hot path.
 * TEST INFRASTRUCTURE ONLY (see tk_oracle.h for the parity status of each
 * function).  Build: oracle/Makefile, flags pinned to
 *   -O2 -mavx -ffp-contract=off
 * so that no fp32 multiply-add is contracted and every operation rounds once.
 * Citations are file:line in the reference tree.
 */
#include "tk_oracle.h"
/* The ONLY thing the oracle shares with the product is the software exponential (tk_dm_expf / tk_dm_sigmoidf):
 * glibc's expf and the GPU's differ by ulps, so a bit-exact score contract needs one source for it
 * (tests/test_post_crosscheck.py holds it against libm within 2 ulp).  Float->int truncation, NaN canonicalisation and
 * the score order are stated here on their own (or_* below), not taken from that header. */
#include "../include/tk_b200_detmath.h"

#include <limits.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

/* (int)x of src/vision/tk_object_detector.c:347-350 where the C cast is defined; saturating outside int range and
 * 0 for NaN, where it is not (the defined behaviour this build fixes on both sides). */
static int32_t or_trunc_i32(float x) {
    if (isnan(x)) return 0;
    const double d = trunc((double)x);
    if (d >= 2147483647.0) return INT32_MAX;
    if (d <= -2147483648.0) return INT32_MIN;
    return (int32_t)d;
}
/* an arithmetic NaN has no portable payload: one quiet NaN pattern for all of them */
static float or_canon_nan(float x) {
    if (!isnan(x)) return x;
    const uint32_t q = 0x7fc00000u;
    float f; memcpy(&f, &q, 4); return f;
}
/* IEEE-754 totalOrder of two floats as -1 / 0 / +1 (sign-magnitude integers: -NaN < -inf < ... < -0 < +0 < ... < +NaN) */
static int or_total_order(float a, float b) {
    int32_t ia, ib;
    memcpy(&ia, &a, 4); memcpy(&ib, &b, 4);
    const int64_t ka = ia < 0 ? -(int64_t)(ia & 0x7fffffff) - 1 : (int64_t)ia;
    const int64_t kb = ib < 0 ? -(int64_t)(ib & 0x7fffffff) - 1 : (int64_t)ib;
    return ka < kb ? -1 : (ka > kb ? 1 : 0);
}

/* ------------------------------------------------------------------------- */
/* Letterbox geometry, integer-only so host launcher and oracle cannot differ. */
/* Unpinned: the reference resizes by plain stretch                            */
/* (src/vision/tk_image_preprocessor.c:43-53).                                 */
/* ------------------------------------------------------------------------- */
static void letterbox_geometry(uint32_t sw, uint32_t sh, uint32_t dw, uint32_t dh,
                               uint32_t* nw, uint32_t* nh, uint32_t* pl, uint32_t* pt) {
    uint64_t a = (uint64_t)dw * sh, b = (uint64_t)dh * sw;
    if (a <= b) {            /* width-limited: scale = dw/sw */
        *nw = dw;
        uint64_t v = (2ull * sh * dw + sw) / (2ull * sw);
        *nh = (uint32_t)(v < 1 ? 1 : (v > dh ? dh : v));
    } else {
        *nh = dh;
        uint64_t v = (2ull * sw * dh + sh) / (2ull * sh);
        *nw = (uint32_t)(v < 1 ? 1 : (v > dw ? dw : v));
    }
    *pl = (dw - *nw) / 2;
    *pt = (dh - *nh) / 2;
}

/* (a2) src/vision/tk_image_preprocessor.c:21-166.
 * ratio :43-44, gather/weights :46-69, plane order :72-74, AVX normalise
 * :84-113 ((v*(1/255)-mean)/std), scalar normalise :156-160 ((v/255-mean)/std).
 * The reference indexes (y*width+x)*3 (stride ignored, :58-61); with
 * src_stride == src_w*3 this is identical. */
int oracle_preprocess_rgb8(const uint8_t* src, uint32_t src_w, uint32_t src_h,
                           uint32_t src_stride, float* out, uint32_t dst_w,
                           uint32_t dst_h, const float mean[3], const float std_dev[3],
                           int norm_variant, int resize_mode, float pad_value) {
    if (!src || !out || dst_w == 0 || dst_h == 0 || src_w < 2 || src_h < 2) return -1;
    uint32_t nw = dst_w, nh = dst_h, pl = 0, pt = 0;
    if (resize_mode == ORACLE_RESIZE_LETTERBOX)
        letterbox_geometry(src_w, src_h, dst_w, dst_h, &nw, &nh, &pl, &pt);
    const size_t npix = (size_t)dst_w * dst_h;
    const float x_ratio = ((float)src_w - 1.0f) / (float)nw;
    const float y_ratio = ((float)src_h - 1.0f) / (float)nh;
    const float inv255 = 1.0f / 255.0f;
    for (uint32_t h = 0; h < dst_h; ++h) {
        for (uint32_t w = 0; w < dst_w; ++w) {
            float v[3];
            if (h < pt || h >= pt + nh || w < pl || w >= pl + nw) {
                v[0] = v[1] = v[2] = pad_value;
            } else {
                float gx = x_ratio * (float)(w - pl);
                float gy = y_ratio * (float)(h - pt);
                int x = (int)gx, y = (int)gy;
                float xd = gx - (float)x, yd = gy - (float)y;
                const uint8_t* r0 = src + (size_t)y * src_stride + (size_t)x * 3;
                const uint8_t* r1 = r0 + src_stride;
                for (int c = 0; c < 3; ++c) {
                    float p1 = (float)r0[c], p2 = (float)r0[3 + c];
                    float p3 = (float)r1[c], p4 = (float)r1[3 + c];
                    v[c] = p1 * (1 - xd) * (1 - yd) + p2 * xd * (1 - yd) +
                           p3 * (1 - xd) * yd + p4 * xd * yd;
                }
            }
            for (int c = 0; c < 3; ++c) {
                float t = (norm_variant == ORACLE_NORM_SCALAR) ? v[c] / 255.0f : v[c] * inv255;
                out[c * npix + (size_t)h * dst_w + w] = (t - mean[c]) / std_dev[c];
            }
        }
    }
    return 0;
}

/* Unpinned (no YUV arithmetic exists in the reference).  Plane layout:
 * src/ffi/c_api/tk_ffi_api.c:917-931.  BT.601 full range in 16.16 fixed point. */
static inline uint8_t clamp_u8(int v) { return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v)); }
void oracle_i420_to_rgb8(const uint8_t* i420, uint32_t w, uint32_t h, uint8_t* rgb) {
    const uint8_t* Y = i420;
    const uint8_t* U = Y + (size_t)w * h;
    const uint8_t* V = U + (size_t)(w / 2) * (h / 2);
    const uint32_t cw = w / 2, ch = h / 2;
    for (uint32_t y = 0; y < h; ++y)
        for (uint32_t x = 0; x < w; ++x) {
            uint32_t cy = y / 2, cx = x / 2;
            if (cy >= ch) cy = ch - 1;      /* odd dimensions: replicate last chroma */
            if (cx >= cw) cx = cw - 1;
            int yy = Y[(size_t)y * w + x];
            int u = (int)U[(size_t)cy * cw + cx] - 128;
            int v = (int)V[(size_t)cy * cw + cx] - 128;
            int r = (yy * 65536 + 91881 * v + 32768) >> 16;
            int g = (yy * 65536 - 22554 * u - 46802 * v + 32768) >> 16;
            int b = (yy * 65536 + 116130 * u + 32768) >> 16;
            uint8_t* o = rgb + ((size_t)y * w + x) * 3;
            o[0] = clamp_u8(r); o[1] = clamp_u8(g); o[2] = clamp_u8(b);
        }
}

uint16_t oracle_f32_to_f16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t mant = x & 0x007fffffu;
    int32_t exp = (int32_t)((x >> 23) & 0xff);
    if (exp == 0xff) return (uint16_t)(sign | 0x7c00u | (mant ? (0x200u | (mant >> 13)) : 0));
    int32_t e = exp - 127 + 15;
    if (e >= 31) return (uint16_t)(sign | 0x7c00u);
    if (e <= 0) {
        if (e < -10) return (uint16_t)sign;
        mant |= 0x00800000u;
        uint32_t shift = (uint32_t)(14 - e);
        uint32_t hm = mant >> shift;
        uint32_t rem = mant & ((1u << shift) - 1u), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (hm & 1u))) hm++;
        return (uint16_t)(sign | hm);
    }
    uint32_t hv = ((uint32_t)e << 10) | (mant >> 13);
    uint32_t rem = mant & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (hv & 1u))) hv++;   /* may carry into inf: correct */
    return (uint16_t)(sign | hv);
}
void oracle_f32_to_f16_array(const float* in, uint16_t* out, size_t n) {
    for (size_t i = 0; i < n; ++i) out[i] = oracle_f32_to_f16(in[i]);
}

float oracle_det_expf(float x) { return tk_dm_expf(x); }
float oracle_det_sigmoidf(float x) { return tk_dm_sigmoidf(x); }

/* (a5) src/vision/tk_object_detector.c:343-355.
 * Pinned fragments: row pointer :344, strict `row[4] > thr` with no activation
 * :345, box truncation :347-350.  Fixed by this build (":351 ... find best class
 * and confidence ..." is elided in the reference): class = first max over the
 * raw class logits (tie rule of src/vision/tk_attribute_classifier.c:83-88),
 * confidence = act(obj) * act(cls). */
uint32_t oracle_decode(const float* logits, uint32_t P, uint32_t C, float thr,
                       int act, oracle_candidate_t* out, uint32_t cap) {
    const size_t L = 5 + (size_t)C;
    uint32_t total = 0;
    for (uint32_t i = 0; i < P; ++i) {
        const float* row = logits + (size_t)i * L;
        float obj = (act == ORACLE_ACT_RAW) ? row[4] : tk_dm_sigmoidf(row[4]);
        if (!(obj > thr)) continue;
        if (total < cap) {
            oracle_candidate_t* c = &out[total];
            c->index = (int32_t)i;
            c->x = or_trunc_i32(row[0] - row[2] / 2);
            c->y = or_trunc_i32(row[1] - row[3] / 2);
            c->w = or_trunc_i32(row[2]);
            c->h = or_trunc_i32(row[3]);
            uint32_t best = 0;
            for (uint32_t j = 1; j < C; ++j)
                if (row[5 + j] > row[5 + best]) best = j;
            float cls;
            if (C == 0) {
                cls = 1.0f;
            } else if (act == ORACLE_ACT_RAW) {
                cls = row[5 + best];
            } else if (act == ORACLE_ACT_SIGMOID) {
                cls = tk_dm_sigmoidf(row[5 + best]);
            } else {   /* class softmax, arithmetic order of tests/tk_gpu_softmax_test.cpp:31-49 */
                float mx = row[5 + best];
                float sum = 0.0f;
                for (uint32_t j = 0; j < C; ++j) sum += tk_dm_expf(row[5 + j] - mx);
                cls = tk_dm_expf(row[5 + best] - mx) / sum;
            }
            c->class_id = best;
            c->confidence = or_canon_nan(obj * cls);
        }
        ++total;
    }
    return total;
}

/* (a6) src/vision/src/object_analysis.rs:71-91 on the truncated integer rects
 * (i32 arithmetic wraps as release-mode Rust does). */
static inline int32_t wadd(int32_t a, int32_t b) { return (int32_t)((uint32_t)a + (uint32_t)b); }
static inline int32_t wsub(int32_t a, int32_t b) { return (int32_t)((uint32_t)a - (uint32_t)b); }
static inline int32_t wmul(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }
float oracle_iou(const oracle_candidate_t* a, const oracle_candidate_t* b) {
    int32_t x_left = a->x > b->x ? a->x : b->x;
    int32_t y_top = a->y > b->y ? a->y : b->y;
    int32_t ar = wadd(a->x, a->w), br = wadd(b->x, b->w);
    int32_t ab = wadd(a->y, a->h), bb = wadd(b->y, b->h);
    int32_t x_right = ar < br ? ar : br;
    int32_t y_bottom = ab < bb ? ab : bb;
    if (x_right < x_left || y_bottom < y_top) return 0.0f;
    float inter = (float)wsub(x_right, x_left) * (float)wsub(y_bottom, y_top);
    float a1 = (float)wmul(a->w, a->h);
    float a2 = (float)wmul(b->w, b->h);
    float uni = a1 + a2 - inter;
    return uni > 0.0f ? inter / uni : 0.0f;
}

/* (a7) greedy, sequential, class-wise.  Order and predicate follow
 * src/gpu/metal/tk_metal_kernels.metal:864-875 (descending score, a box is
 * dropped when IoU with an already-kept box is strictly > thr), restricted to
 * equal class_id (north star) and evaluated sequentially.  Ties: proposal index
 * ascending.  Scores are ordered by IEEE-754 totalOrder (or_total_order). */
static int cand_cmp(const void* pa, const void* pb) {
    const oracle_candidate_t* a = (const oracle_candidate_t*)pa;
    const oracle_candidate_t* b = (const oracle_candidate_t*)pb;
    const int o = or_total_order(a->confidence, b->confidence);
    if (o != 0) return -o;                               /* higher score first */
    if (a->index != b->index) return a->index < b->index ? -1 : 1;
    return 0;
}
uint32_t oracle_nms_classwise(const oracle_candidate_t* cands, uint32_t count, float iou_thr,
                              uint32_t max_det, oracle_candidate_t* kept) {
    if (count == 0 || max_det == 0) return 0;
    oracle_candidate_t* s = (oracle_candidate_t*)malloc((size_t)count * sizeof(*s));
    uint32_t* kidx = (uint32_t*)malloc((size_t)count * sizeof(uint32_t));
    memcpy(s, cands, (size_t)count * sizeof(*s));
    qsort(s, count, sizeof(*s), cand_cmp);
    uint32_t nk = 0;
    for (uint32_t k = 0; k < count; ++k) {
        int drop = 0;
        for (uint32_t q = 0; q < nk && !drop; ++q) {
            const oracle_candidate_t* j = &s[kidx[q]];
            if (j->class_id == s[k].class_id && oracle_iou(j, &s[k]) > iou_thr) drop = 1;
        }
        if (!drop) kidx[nk++] = k;
    }
    uint32_t n = nk < max_det ? nk : max_det;
    for (uint32_t q = 0; q < n; ++q) kept[q] = s[kidx[q]];
    free(s); free(kidx);
    return n;
}

/* (a9) src/vision/tk_object_detector.c:372-375 is empty; unpinned.  Stretch
 * inverse in fp32 with truncation, matching the (int) convention of :347-350. */
void oracle_restore_coords(oracle_candidate_t* d, uint32_t n, uint32_t src_w, uint32_t src_h,
                           uint32_t in_w, uint32_t in_h) {
    float sx = (float)src_w / (float)in_w, sy = (float)src_h / (float)in_h;
    for (uint32_t i = 0; i < n; ++i) {
        d[i].x = or_trunc_i32((float)d[i].x * sx);
        d[i].y = or_trunc_i32((float)d[i].y * sy);
        d[i].w = or_trunc_i32((float)d[i].w * sx);
        d[i].h = or_trunc_i32((float)d[i].h * sy);
    }
}

/* (a10) tests/tk_gpu_softmax_test.cpp:24-51: max :33-38, in-order sum :41-44,
 * divide :47-49.  use_det_exp=0 uses libm expf like the reference's std::exp. */
void oracle_softmax_rows(const float* in, float* out, uint32_t rows, uint32_t cols, int det) {
    for (uint32_t i = 0; i < rows; ++i) {
        const float* r = in + (size_t)i * cols;
        float* o = out + (size_t)i * cols;
        float mx = r[0];
        for (uint32_t j = 1; j < cols; ++j) if (r[j] > mx) mx = r[j];
        float sum = 0.0f;
        for (uint32_t j = 0; j < cols; ++j) sum += det ? tk_dm_expf(r[j] - mx) : expf(r[j] - mx);
        for (uint32_t j = 0; j < cols; ++j) o[j] = (det ? tk_dm_expf(r[j] - mx) : expf(r[j] - mx)) / sum;
    }
}
void oracle_argmax_rows(const float* in, uint32_t* out, uint32_t rows, uint32_t cols) {
    for (uint32_t i = 0; i < rows; ++i) {
        const float* r = in + (size_t)i * cols;
        uint32_t best = 0;
        for (uint32_t j = 1; j < cols; ++j) if (r[j] > r[best]) best = j;
        out[i] = best;
    }
}

/* (a11) src/vision/tk_attribute_classifier.c:19-36 (rgb_to_hsv) and :54-88.
 * The hue arithmetic mixes float and double exactly as the C source does:
 * `2.0 + (b-r)/delta`, `*h *= 60.0`, `*h += 360.0` are evaluated in double and
 * rounded to float on assignment; `s < 0.1`, `v < 0.1`, `v > 0.9` compare the
 * float promoted to double; `h < 30` etc. compare in float. */
static int color_bin(uint8_t R, uint8_t G, uint8_t B) {
    float r = R / 255.0f, g = G / 255.0f, b = B / 255.0f;
    float mx = fmaxf(r, fmaxf(g, b)), mn = fminf(r, fminf(g, b));
    float delta = mx - mn, v = mx, s, h = 0.0f;
    int gray;
    if (mx > 0.0) {
        s = delta / mx;
        if (r >= mx) h = (g - b) / delta;
        else if (g >= mx) h = (float)(2.0 + (double)((b - r) / delta));
        else h = (float)(4.0 + (double)((r - g) / delta));
        h = (float)((double)h * 60.0);
        if ((double)h < 0.0) h = (float)((double)h + 360.0);
    } else {
        s = 0.0f;
    }
    gray = (double)s < 0.1;
    if (gray) {
        if ((double)v < 0.1) return 6;
        if ((double)v > 0.9) return 7;
        return 8;
    }
    if (h < 30 || h >= 330) return 0;
    if (h >= 30 && h < 90) return 1;
    if (h >= 90 && h < 150) return 2;
    if (h >= 150 && h < 210) return 3;
    if (h >= 210 && h < 270) return 4;
    if (h >= 270 && h < 330) return 5;
    return -1;   /* NaN hue with s >= 0.1 cannot occur; counted nowhere, as in the reference */
}
int oracle_dominant_color(const uint8_t* rgb, uint32_t w, uint32_t h, uint32_t stride,
                          int bx, int by, int bw, int bh, int64_t bins[9]) {
    for (int i = 0; i < 9; ++i) bins[i] = 0;
    for (int y = by; y < by + bh; ++y)
        for (int x = bx; x < bx + bw; ++x) {
            if (x < 0 || (uint32_t)x >= w || y < 0 || (uint32_t)y >= h) continue;
            const uint8_t* p = rgb + (size_t)y * stride + (size_t)x * 3;
            int k = color_bin(p[0], p[1], p[2]);
            if (k >= 0) bins[k]++;
        }
    int best = 0;
    for (int i = 1; i < 9; ++i) if (bins[i] > bins[best]) best = i;
    return best;
}

/* src/vision/tk_attribute_classifier.c:119-147: integer luminance
 * (299R+587G+114B)/1000 truncated to u8, |above-below| > 100, density
 * count/(w*h) > 0.1 (float promoted to double).  The reference does not clip
 * the bbox to the frame; callers must pass an in-frame bbox. */
static inline int gray_u8(const uint8_t* p) { return (uint8_t)((p[0] * 299 + p[1] * 587 + p[2] * 114) / 1000); }
int oracle_door_state(const uint8_t* rgb, uint32_t w, uint32_t h, uint32_t stride,
                      int bx, int by, int bw, int bh, int32_t* edge_pixels) {
    (void)w; (void)h;
    int count = 0;
    for (int y = by + 1; y < by + bh - 1; ++y)
        for (int x = bx + 1; x < bx + bw - 1; ++x) {
            const uint8_t* above = rgb + (size_t)(y - 1) * stride + (size_t)x * 3;
            const uint8_t* below = rgb + (size_t)(y + 1) * stride + (size_t)x * 3;
            if (abs(gray_u8(above) - gray_u8(below)) > 100) count++;
        }
    if (edge_pixels) *edge_pixels = count;
    float density = (float)count / (float)(bw * bh);
    return (double)density > 0.1 ? 1 : 0;
}

/* ---- depth post-processing --------------------------- */

/* src/vision/tk_depth_midas.c:473-499 (convert_inverse_depth_to_metric),
 * statement for statement: extrema seeded with data[0] and updated with strict
 * compares (a NaN never replaces them; a NaN seed sticks), `max - min < 1e-6f`
 * -> every element = max_depth, else max_depth - (d-min)/(max-min)*(max_depth-min_depth).
 * Built with -ffp-contract=off: one rounding per operation.  PARITY UNPINNED BY
 * EXECUTION — the reference file cannot compile here (ONNX Runtime C++ API inside
 * a .c file) and no reference test pins it; the restatement is the code above
 * re-typed. */
void oracle_inverse_depth_to_metric(const float* in, float* out, size_t n, float min_depth, float max_depth) {
    if (n == 0) return;
    float min_val = in[0], max_val = in[0];
    for (size_t i = 1; i < n; i++) {
        if (in[i] < min_val) min_val = in[i];
        if (in[i] > max_val) max_val = in[i];
    }
    if (max_val - min_val < 1e-6f) {
        for (size_t i = 0; i < n; i++) out[i] = max_depth;
        return;
    }
    for (size_t i = 0; i < n; i++) {
        float normalized = (in[i] - min_val) / (max_val - min_val);
        out[i] = max_depth - normalized * (max_depth - min_depth);
    }
}

/* Rust `f32 as u32`: saturating, NaN -> 0. */
static uint32_t rust_f32_as_u32(float f) {
    if (!(f > 0.0f)) return 0;                /* negatives, -0, NaN */
    if (f >= 4294967296.0f) return 0xffffffffu;
    return (uint32_t)f;
}
static int cmp_f32(const void* a, const void* b) {
    float x = *(const float*)a, y = *(const float*)b;
    return (x > y) - (x < y);
}
/* src/vision/src/object_analysis.rs:227-279 (calculate_raw_distance): bbox ->
 * normalised corners (f32 divide) -> depth-map pixel via `.round() as u32`
 * (half away from zero, saturating) -> inclusive window -> samples with
 * 0.1 < d < 100 -> if >= 10: sort, q1 = v[len/4], q3 = v[len*3/4], fence
 * q1 - 1.5 iqr .. q3 + 1.5 iqr, mean of the samples inside as a sequential f32
 * sum in ascending order / count; otherwise -1.  i32 `x + w` wraps (release
 * build).  `scratch` must hold depth_w*depth_h floats.
 * Pinned by the reference's own unit test (object_analysis.rs:286-347): a
 * 20x20 box over a 100x100 map of 10.0 -> 10.0, of 12.0 -> 12.0 (the Kalman
 * smoothing that follows is downstream, out of scope). */
float oracle_detection_distance(int bx, int by, int bw, int bh, const float* depth, uint32_t depth_w,
                                uint32_t depth_h, uint32_t frame_w, uint32_t frame_h, float* scratch) {
    float fw = (float)frame_w, fh = (float)frame_h;
    int xr = (int)((uint32_t)bx + (uint32_t)bw), yb = (int)((uint32_t)by + (uint32_t)bh);
    float nx0 = (float)bx / fw, ny0 = (float)by / fh, nx1 = (float)xr / fw, ny1 = (float)yb / fh;
    uint32_t x0 = rust_f32_as_u32(roundf(nx0 * (float)(depth_w - 1)));
    uint32_t y0 = rust_f32_as_u32(roundf(ny0 * (float)(depth_h - 1)));
    uint32_t x1 = rust_f32_as_u32(roundf(nx1 * (float)(depth_w - 1)));
    uint32_t y1 = rust_f32_as_u32(roundf(ny1 * (float)(depth_h - 1)));
    if (x0 >= depth_w || y0 >= depth_h || x1 >= depth_w || y1 >= depth_h || x0 >= x1 || y0 >= y1) return -1.0f;
    size_t n = 0;
    for (uint32_t y = y0; y <= y1; ++y)
        for (uint32_t x = x0; x <= x1; ++x) {
            float d = depth[(size_t)y * depth_w + x];
            if (d > 0.1f && d < 100.0f) scratch[n++] = d;
        }
    if (n < 10) return -1.0f;
    qsort(scratch, n, sizeof(float), cmp_f32);
    float q1 = scratch[n / 4], q3 = scratch[n * 3 / 4];
    float iqr = q3 - q1;
    float lower = q1 - 1.5f * iqr, upper = q3 + 1.5f * iqr;
    float sum = 0.0f;
    size_t cnt = 0;
    for (size_t i = 0; i < n; ++i)
        if (scratch[i] >= lower && scratch[i] <= upper) { sum += scratch[i]; cnt++; }
    return cnt ? sum / (float)cnt : -1.0f;
}
