// exp_ffma2.cu — issue rate of packed fp32 (FFMA2 / FMUL2) against scalar FFMA and against the byte-permute stream the
// pre-process kernel mixes it with.  One CTA of 1024 threads per SM (8 warps per scheduler), eight independent chains per
// thread, cycles from clock64 around the loop: warp instructions per cycle per scheduler.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/exp_ffma2 tools/exp_ffma2.cu && /tmp/exp_ffma2
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITER 2048

#define F2(op, d, a, b, c) asm volatile(op " %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c))
#define M2(op, d, a, b) asm volatile(op " %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b))

template <int MODE>
__global__ void __launch_bounds__(1024, 1) k(unsigned long long* out, long long* cyc, float fone, unsigned long long one2, uint32_t sel) {
    unsigned long long a[8], x = one2 + threadIdx.x, y = one2;
    float s[8], fx = fone + threadIdx.x, fy = fone;
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = one2 + i; s[i] = fone + i; w[i] = threadIdx.x * 9u + i; }
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < ITER; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s[i]) : "f"(fx), "f"(fy));
            if (MODE == 1) F2("fma.rn.f32x2", a[i], a[i], x, y);
            if (MODE == 2) M2("mul.rn.f32x2", a[i], a[i], x);
            if (MODE == 3) M2("add.rn.f32x2", a[i], a[i], x);
            if (MODE == 4) asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel), "r"(sel + i));
            if (MODE == 5) { F2("fma.rn.f32x2", a[i], a[i], x, y); asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel), "r"(sel + i)); }
            if (MODE == 6) { asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s[i]) : "f"(fx), "f"(fy)); asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel), "r"(sel + i)); }
            if (MODE == 7) { F2("fma.rn.f32x2", a[i], a[i], x, y); asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel), "r"(sel + i));
                             asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel + 1), "r"(sel + i)); }
            if (MODE == 8) { M2("mul.rn.f32x2", a[i], a[i], x); asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel), "r"(sel + i)); }
            if (MODE == 9) { asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s[i]) : "f"(fx), "f"(fy)); asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(s[(i + 4) & 7]) : "f"(fy), "f"(fx));
                             asm volatile("prmt.b32 %0, %0, %1, %2;" : "+r"(w[i]) : "r"(sel), "r"(sel + i)); }
        }
    }
    const long long t1 = clock64();
    unsigned long long acc = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += a[i] + __float_as_uint(s[i]) + w[i];
    out[blockIdx.x * 1024 + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static void run(const char* name, int per_iter, unsigned long long* out, long long* cyc, int sms) {
    const float one = 1.0f;
    unsigned long long one2; { uint32_t b; memcpy(&b, &one, 4); one2 = ((unsigned long long)b << 32) | b; }
    k<MODE><<<sms, 1024>>>(out, cyc, one, one2, 0x3210u);
    k<MODE><<<sms, 1024>>>(out, cyc, one, one2, 0x3210u);
    cudaDeviceSynchronize();
    long long h[256]; cudaMemcpy(h, cyc, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < sms; ++i) avg += (double)h[i]; avg /= sms;
    // warp instructions issued per scheduler: 8 warps x ITER x 8 x per_iter
    const double inst = 8.0 * ITER * 8.0 * per_iter;
    printf("%-34s %8.0f cycles  %.3f warp-inst/cycle/scheduler  (%.2f cycles per instruction)\n", name, avg, inst / avg, avg / inst);
}

int main() {
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    unsigned long long* out; long long* cyc;
    cudaMalloc(&out, sizeof(unsigned long long) * 1024 * sms); cudaMalloc(&cyc, sizeof(long long) * sms);
    run<0>("FFMA  (3 registers)", 1, out, cyc, sms);
    run<1>("FFMA2 (3 register pairs)", 1, out, cyc, sms);
    run<2>("FMUL2", 1, out, cyc, sms);
    run<3>("FADD2", 1, out, cyc, sms);
    run<4>("PRMT", 1, out, cyc, sms);
    run<5>("FFMA2 + PRMT", 2, out, cyc, sms);
    run<6>("FFMA + PRMT", 2, out, cyc, sms);
    run<7>("FFMA2 + 2 PRMT", 3, out, cyc, sms);
    run<8>("FMUL2 + PRMT", 2, out, cyc, sms);
    run<9>("2 FFMA + PRMT", 3, out, cyc, sms);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
