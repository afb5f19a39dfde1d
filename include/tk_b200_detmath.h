/*
 * tk_b200_detmath.h — deterministic scalar math shared by host and device.
 *
 * glibc's expf and CUDA's expf differ by ulps, which is enough to flip a strict
 * `>` threshold or a sort order.  Everything here is built only from IEEE-754
 * correctly rounded primitives (add, mul, fma, divide, float<->int conversion),
 * so the same source gives the same bits under gcc (-ffp-contract=off, fmaf from
 * libm or hardware) and under nvcc (fmaf -> FFMA, __fdiv_rn).
 *
 * Reference context: the sigmoid the reference's CUDA side defines is
 * 1/(1+exp(-x)) (src/gpu/extensions/cuda/tk_cuda_tensor_ops.cu:181,
 * src/gpu/cuda/tk_cuda_math_helpers.cu:21-27) and its softmax is
 * exp(x-max)/sum (tests/tk_gpu_softmax_test.cpp:24-51); neither pins an exp
 * implementation.
 */
#ifndef TK_B200_DETMATH_H
#define TK_B200_DETMATH_H

#include <stdint.h>
#include <string.h>
#ifndef __CUDACC__
#include <math.h>
#endif

#ifdef __CUDACC__
#define TK_DM_FN __host__ __device__ __forceinline__
#else
#define TK_DM_FN static inline
#endif

TK_DM_FN float tk_dm_bits_to_float(uint32_t u) {
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}
TK_DM_FN uint32_t tk_dm_float_to_bits(float f) {
#ifdef __CUDA_ARCH__
    return __float_as_uint(f);
#else
    uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
TK_DM_FN float tk_dm_mul(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
}
TK_DM_FN float tk_dm_add(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
}
TK_DM_FN float tk_dm_sub(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fsub_rn(a, b);
#else
    return a - b;
#endif
}
TK_DM_FN float tk_dm_div(float a, float b) {
#ifdef __CUDA_ARCH__
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
}

/* exp(x) in fp32, |err| < 3 ulp (measured 2.62), bit-identical on host and device.
 * n = round(x*log2e); r = x - n*ln2 (two-step Cody-Waite with fma);
 * e^r by a degree-6 Horner polynomial in fma; result scaled by 2^n through two
 * exact power-of-two multiplies so that gradual underflow is handled by the
 * hardware multiply identically on both sides. */
TK_DM_FN float tk_dm_expf(float x) {
    if (x != x) return x;
    if (x > 88.72283935546875f) return tk_dm_bits_to_float(0x7f800000u);
    if (x < -103.97208404541015625f) return 0.0f;
    const float log2e  = 1.44269502162933349609375f;
    const float ln2_hi = 0.693145751953125f;           /* 12 significant bits */
    const float ln2_lo = 1.42860676533018704503775e-06f;
    float t = tk_dm_mul(x, log2e);
    /* round-to-nearest-even via the 1.5*2^23 trick (|t| < 2^22 here) */
    const float magic = 12582912.0f;
    float nf = tk_dm_sub(tk_dm_add(t, magic), magic);
    float r = fmaf(-nf, ln2_hi, x);
    r = fmaf(-nf, ln2_lo, r);
    float p = 1.0f / 720.0f;
    p = fmaf(p, r, 1.0f / 120.0f);
    p = fmaf(p, r, 1.0f / 24.0f);
    p = fmaf(p, r, 1.0f / 6.0f);
    p = fmaf(p, r, 0.5f);
    p = fmaf(p, r, 1.0f);
    p = fmaf(p, r, 1.0f);
    int n = (int)nf;                       /* exact: nf is integral, |nf| <= 150 */
    int n1 = n / 2, n2 = n - n1;           /* C truncation on both sides         */
    float s1 = tk_dm_bits_to_float((uint32_t)(n1 + 127) << 23);
    float s2 = tk_dm_bits_to_float((uint32_t)(n2 + 127) << 23);
    return tk_dm_mul(tk_dm_mul(p, s1), s2);
}

/* 1/(1+exp(-x)) with a correctly rounded divide. */
TK_DM_FN float tk_dm_sigmoidf(float x) {
    return tk_dm_div(1.0f, tk_dm_add(1.0f, tk_dm_expf(-x)));
}

/* Saturating float->int truncation with CUDA's cvt.rzi.s32.f32 semantics
 * (NaN -> 0).  Equals the C cast `(int)x` of
 * src/vision/tk_object_detector.c:347-350 for every in-range x; out of range
 * the C cast is undefined (x86 gives INT_MIN), so the saturating form is the
 * defined behaviour of this build on both sides. */
TK_DM_FN int32_t tk_dm_f2i_trunc(float x) {
#ifdef __CUDA_ARCH__
    return __float2int_rz(x);
#else
    if (x != x) return 0;
    if (x >= 2147483648.0f) return INT32_MAX;
    if (x <= -2147483648.0f) return INT32_MIN;
    return (int32_t)x;
#endif
}

/* NaN results are canonicalised to the quiet NaN 0x7fc00000: x86 SSE propagates
 * the operand's payload while the GPU's FMUL returns 0x7fffffff, so an
 * arithmetic NaN would otherwise differ in its bits between host and device. */
TK_DM_FN float tk_dm_canon_nan(float x) {
    return (x != x) ? tk_dm_bits_to_float(0x7fc00000u) : x;
}

/* Order-preserving map float -> u32 (total order on IEEE bit patterns:
 * -NaN < -inf < ... < -0 < +0 < ... < +inf < +NaN). */
TK_DM_FN uint32_t tk_dm_float_order_key(float f) {
    uint32_t b = tk_dm_float_to_bits(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

#endif /* TK_B200_DETMATH_H */
