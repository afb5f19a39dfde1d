// Experiment: decode pass 1 (one 4-byte objectness word out of every 340-byte row) as a TMA GATHER.
// The logits [rows][85] f32 are viewed as [rows/4][340] f32 (row pitch 1360 B = 85 * 16: a legal TMA stride, which 340 B
// is not); the objectness words of the four rows of a super-row sit at floats 4, 89, 174, 259.  One
// cp.async.bulk.tensor.2d with a box of (4 floats, R super-rows) pulls 16 bytes around each word for R rows in ONE
// instruction — no LSU request per row, no register per load.  Question: does the DRAM side get cheaper than the
// 64 B / row the LSU path fetches (tools/exp_stride.cu: 25 us for 64 x 25200 rows)?
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/exp_stride_tma tools/exp_stride_tma.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int R>     // super-rows per box
__global__ void __launch_bounds__(256) scan_tma(const __grid_constant__ CUtensorMap map, uint32_t super_rows, unsigned* out, int kmask) {
    __shared__ __align__(128) float buf[2][4][R][4];      // [stage][word k][super-row][4 floats]
    __shared__ __align__(8) uint64_t bar[2];
    const uint32_t tid = threadIdx.x;
    if (tid == 0) {
        for (int s = 0; s < 2; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[s])));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    const uint32_t chunks = (super_rows + R - 1) / R;
    unsigned c = 0;
    auto issue = [&](uint32_t ch, int s) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[s])), "r"((uint32_t)__popc(kmask) * R * 16u) : "memory");
        for (int k = 0; k < 4; ++k)
            if ((kmask >> k) & 1)
            asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                         ::"r"(smem_u32(&buf[s][k][0][0])), "l"(&map), "r"(4 + 85 * k), "r"((int)(ch * R)), "r"(smem_u32(&bar[s])) : "memory");
    };
    uint32_t it = 0;
    uint32_t ch = blockIdx.x;
    if (tid == 0 && ch < chunks) issue(ch, 0);
    for (; ch < chunks; ch += gridDim.x, ++it) {
        const int s = it & 1;
        const uint32_t nxt = ch + gridDim.x;
        if (tid == 0 && nxt < chunks) issue(nxt, s ^ 1);
        uint32_t ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar[s])), "r"((it >> 1) & 1u) : "memory");
        for (uint32_t i = tid; i < 4u * R; i += 256) c += buf[s][i / R][i % R][0] > 0.5f;
        __syncthreads();      // everyone is done with stage s before it is refilled two iterations later
    }
    if (c == 0xffffffffu) *out = c;
}

int main(int argc, char** argv) {
    const int kmask = argc > 1 ? atoi(argv[1]) : 15;
    const size_t rows = 64ull * 25200;
    const int L = 85;
    float* x; unsigned* out;
    cudaMalloc(&x, rows * L * 4 + 4096); cudaMalloc(&out, 4);
    cudaMemset(x, 0, rows * L * 4);
    EncodeFn encode = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &q) != cudaSuccess || !encode) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    const int promos[3] = {CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B};
    const char* pn[3] = {"L2 promotion NONE", "L2 promotion 64B", "L2 promotion 128B"};
    for (int pi = 0; pi < 3; ++pi) {
        constexpr int R = 256;
        CUtensorMap map;
        cuuint64_t dims[2] = {340, rows / 4};
        cuuint64_t strides[1] = {1360};
        cuuint32_t box[2] = {4, R};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, x, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, (CUtensorMapL2promotion)promos[pi], CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
        for (int grid : {148, 296, 592}) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            for (int w = 0; w < 3; ++w) scan_tma<R><<<grid, 256>>>(map, (uint32_t)(rows / 4), out, kmask);
            cudaEventRecord(e0);
            for (int w = 0; w < 20; ++w) scan_tma<R><<<grid, 256>>>(map, (uint32_t)(rows / 4), out, kmask);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            const double us = ms * 1000 / 20;
            printf("kmask %d TMA gather, box 4 x %d, %-18s grid %3d: %7.1f us -> %5.1f B/row at 6.5 TB/s-equivalent (%s)\n", kmask, R, pn[pi], grid, us,
                   us * 6.5e6 / rows, cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
