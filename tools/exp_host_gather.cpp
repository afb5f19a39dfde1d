// Experiment: how fast can host threads gather the objectness word (4 bytes every 340) of 64 x 25200 rows
// into a dense array?  (Alternative to reading the rows in place over PCIe, one request per row.)
// g++ -O2 -pthread tools/exp_host_gather.cpp -o /tmp/exp_host_gather && /tmp/exp_host_gather
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
int main() {
    const size_t frames = 64, P = 25200, L = 85, rows = frames * P;
    float* src = (float*)aligned_alloc(4096, rows * L * 4);
    float* dst = (float*)aligned_alloc(4096, rows * 4);
    for (size_t i = 0; i < rows * L; i += 16) src[i] = (float)i;
    memset(dst, 0, rows * 4);
    printf("hardware threads: %u\n", std::thread::hardware_concurrency());
    for (int T : {1, 2, 4, 8, 16, 32}) {
        double best = 1e9;
        for (int rep = 0; rep < 5; ++rep) {
            auto t0 = std::chrono::steady_clock::now();
            std::vector<std::thread> th;
            for (int t = 0; t < T; ++t)
                th.emplace_back([=] {
                    size_t a = rows * t / T, b = rows * (t + 1) / T;
                    const float* s = src + a * L + 4;
                    for (size_t r = a; r < b; ++r, s += L) dst[r] = *s;
                });
            for (auto& x : th) x.join();
            double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            if (ms < best) best = ms;
        }
        printf("threads %2d: %.2f ms  (%.1f M rows/s)\n", T, best, rows / best / 1e3);
    }
    return 0;
}
