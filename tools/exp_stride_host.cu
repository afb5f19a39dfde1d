// Experiment: PCIe time of a 4-byte read every 340 bytes of PINNED HOST memory (decode pass 1 reading the logits in place)
// under different load flavours and L2 fetch-granularity limits.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstring>
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
    else if (MODE == 9) asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
    else asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
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
    printf("%-44s L=%d: %8.1f us  -> %6.1f B/row at 55 GB/s-equivalent  (%s)\n", name, L, us, us * 55e3 / rows,
           cudaGetErrorString(cudaGetLastError()));
}
int main() {
    const size_t rows = 64ull * 25200;
    const int L = 85;
    float* h; float* x; unsigned* out;
    cudaHostAlloc(&h, rows * L * 4 + 1024, cudaHostAllocMapped | cudaHostAllocPortable);
    memset(h, 0, rows * L * 4);
    cudaHostGetDevicePointer(&x, h, 0);
    cudaMalloc(&out, 4);
    {   // copy-engine gather: 2-D copy of the objectness column only
        float* d; cudaMalloc(&d, rows * 64);
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (size_t wbytes : {4, 16, 32, 64}) {
            for (int it = 0; it < 2; ++it) {
                cudaEventRecord(e0);
                cudaMemcpy2DAsync(d, wbytes, (const char*)h + 16, (size_t)L * 4, wbytes, rows, cudaMemcpyHostToDevice, 0);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1);
                printf("cudaMemcpy2DAsync width %zu B, pitch 340, %zu rows: %8.1f us (%s)\n", wbytes, rows, ms * 1000, cudaGetErrorString(cudaGetLastError()));
            }
        }
        // plain copy for reference
        cudaEventRecord(e0);
        cudaMemcpyAsync(d, h, rows * 64, cudaMemcpyHostToDevice, 0);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        printf("cudaMemcpyAsync %zu B: %8.1f us = %.1f GB/s\n", rows * 64, ms * 1000, rows * 64 / ms / 1e6);
        cudaFree(d);
    }
    for (size_t g : {128}) {
        cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, g);
        size_t v = 0; cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity);
        printf("--- cudaLimitMaxL2FetchGranularity = %zu\n", v);
        run<0>(x, rows, L, out, "__ldg (ld.global.nc)");
        run<1>(x, rows, L, out, "ld.global.cg");
        run<4>(x, rows, L, out, "ld.volatile.global");
        run<5>(x, rows, L, out, "ld.global.L2::64B");
        run<6>(x, rows, L, out, "ld.global.nc.L1::no_allocate");
        run<7>(x, rows, L, out, "ld.global.nc.L1::no_allocate.L2::64B");
        run<9>(x, rows, L, out, "ld.relaxed.gpu.global");
        run<10>(x, rows, L, out, "ld.relaxed.sys.global");
    }
    return 0;
}
