// exp_upload.cu — how to get one pageable 640x480 RGB frame (921 600 B) into HBM: the driver's pageable
// cudaMemcpyAsync against our own staging through a pinned buffer (CPU memcpy + async copy, 1 / 2 / 4 chunks), with the
// source hot in the CPU caches (a harness that reuses its frame) and cold (a frame fresh from a camera).
//   nvcc -O3 -o tools/_exp_upload tools/exp_upload.cu && tools/_exp_upload
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

static double now_us() { return std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main() {
    const size_t N = 921600;
    uint8_t* src = static_cast<uint8_t*>(malloc(N));
    memset(src, 7, N);
    uint8_t *pin, *dev;
    cudaHostAlloc(reinterpret_cast<void**>(&pin), N, cudaHostAllocDefault);
    cudaMalloc(reinterpret_cast<void**>(&dev), N);
    cudaStream_t st; cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
    const size_t FL = 256u << 20;
    uint8_t* flush = static_cast<uint8_t*>(malloc(FL));
    volatile uint64_t sink = 0;
    for (int cold = 0; cold < 2; ++cold) {
        for (int mode = 0; mode < 4; ++mode) {      // 0: driver pageable copy; 1, 2, 3: own staging in 1, 2, 4 chunks
            const int chunks = mode == 0 ? 0 : (1 << (mode - 1));
            std::vector<double> ts;
            for (int it = 0; it < (cold ? 30 : 200); ++it) {
                if (cold) { memset(flush, it, FL); uint64_t s = 0; for (size_t i = 0; i < FL; i += 64) s += flush[i]; sink += s; }
                const double t0 = now_us();
                if (mode == 0) {
                    cudaMemcpyAsync(dev, src, N, cudaMemcpyHostToDevice, st);
                } else {
                    const size_t cb = (N / chunks + 63) & ~size_t(63);
                    for (size_t o = 0; o < N; o += cb) {
                        const size_t len = std::min(cb, N - o);
                        memcpy(pin + o, src + o, len);
                        cudaMemcpyAsync(dev + o, pin + o, len, cudaMemcpyHostToDevice, st);
                    }
                }
                cudaStreamSynchronize(st);
                ts.push_back(now_us() - t0);
            }
            std::sort(ts.begin(), ts.end());
            printf("%s source, %-34s: median %6.1f us  p10 %6.1f  p90 %6.1f\n", cold ? "cold" : "hot ",
                   mode == 0 ? "cudaMemcpyAsync from pageable" : (chunks == 1 ? "memcpy to pinned + copy, 1 chunk" : (chunks == 2 ? "memcpy to pinned + copy, 2 chunks" : "memcpy to pinned + copy, 4 chunks")),
                   ts[ts.size() / 2], ts[ts.size() / 10], ts[ts.size() * 9 / 10]);
        }
    }
    printf("%s (%llu)\n", cudaGetErrorString(cudaGetLastError()), (unsigned long long)sink);
    return 0;
}
