// Experiment: DRAM traffic/time of a 4-byte read every 340 bytes (decode pass 1) under different load flavours.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
template <int MODE>
__device__ __forceinline__ float ld(const float* p) {
    float v;
    if (MODE == 0) v = __ldg(p);
    else if (MODE == 1) asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 2) asm volatile("ld.global.cs.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 3) asm volatile("ld.global.lu.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 4) asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 5) asm volatile("ld.global.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 6) asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 7) asm volatile("ld.global.nc.L1::no_allocate.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else if (MODE == 8) asm volatile("ld.global.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
}
template <int MODE>
__global__ void scan(const float* __restrict__ x, size_t rows, int L, int off, unsigned* out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t step = (size_t)gridDim.x * blockDim.x;
    unsigned c = 0;
    for (; i + 3 * step < rows; i += 4 * step) {
        float a = ld<MODE>(x + i * L + off), b = ld<MODE>(x + (i + step) * L + off);
        float d = ld<MODE>(x + (i + 2 * step) * L + off), e = ld<MODE>(x + (i + 3 * step) * L + off);
        c += (a > 0.5f) + (b > 0.5f) + (d > 0.5f) + (e > 0.5f);
    }
    for (; i < rows; i += step) c += ld<MODE>(x + i * L + off) > 0.5f;
    if (c == 0xffffffffu) *out = c;
}
template <int MODE>
void run(const float* x, size_t rows, int L, unsigned* out, const char* name) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 3; ++w) scan<MODE><<<148 * 8, 256>>>(x, rows, L, 4, out);
    cudaEventRecord(e0);
    for (int w = 0; w < 20; ++w) scan<MODE><<<148 * 8, 256>>>(x, rows, L, 4, out);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double us = ms * 1000 / 20;
    printf("%-44s L=%d: %7.1f us  -> %6.1f B/row at 6.5 TB/s-equivalent, %6.1f Mrows/s  (%s)\n", name, L, us, us * 6.5e6 / rows, rows / us,
           cudaGetErrorString(cudaGetLastError()));
}
int main(int argc, char** argv) {
    const size_t rows = 64ull * 25200;
    if (argc > 1) {      // ./exp_stride limit: the plain and the hinted loads under each cudaLimitMaxL2FetchGranularity
        for (size_t g : {32, 64, 128}) {
            cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g);
            size_t got = 0; cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
            printf("--- cudaLimitMaxL2FetchGranularity = %zu (%s; in effect %zu)\n", g, cudaGetErrorString(e), got);
            float* x; unsigned* out;
            cudaMalloc(&x, rows * 85 * 4 + 1024); cudaMalloc(&out, 4);
            cudaMemset(x, 0, rows * 85 * 4);
            run<0>(x, rows, 85, out, "__ldg (ld.global.nc)");
            run<6>(x, rows, 85, out, "ld.global.nc.L1::no_allocate");
            run<7>(x, rows, 85, out, "ld.global.nc.L1::no_allocate.L2::64B");
            run<1>(x, rows, 85, out, "ld.global.cg");
            cudaFree(x); cudaFree(out);
        }
        return 0;
    }
    for (int L : {85, 96}) {
        float* x; unsigned* out;
        cudaMalloc(&x, rows * L * 4 + 1024); cudaMalloc(&out, 4);
        cudaMemset(x, 0, rows * L * 4);
        run<0>(x, rows, L, out, "__ldg (ld.global.nc)");
        run<1>(x, rows, L, out, "ld.global.cg");
        run<2>(x, rows, L, out, "ld.global.cs");
        run<3>(x, rows, L, out, "ld.global.lu");
        run<4>(x, rows, L, out, "ld.volatile.global");
        run<5>(x, rows, L, out, "ld.global.L2::64B");
        run<6>(x, rows, L, out, "ld.global.nc.L1::no_allocate");
        run<7>(x, rows, L, out, "ld.global.nc.L1::no_allocate.L2::64B");
        run<8>(x, rows, L, out, "ld.global.L1::no_allocate");
        run<9>(x, rows, L, out, "ld.relaxed.gpu.global");
        cudaFree(x); cudaFree(out);
    }
    return 0;
}
