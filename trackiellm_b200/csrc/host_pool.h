// host_pool.h — a small persistent pool of host threads owned by a handle (pipeline, shard harness).
//
// The reference runs this path from one host thread per device and never parallelises host-side byte movement
// (src/gpu/cuda/tk_cuda_dispatch.cu:163-247 creates three streams and nothing else).  Here the host side of
// tk_b200_pipeline_process_host packs the objectness column of page-locked logits while the copy engine moves the
// frames; round 1 created and joined std::threads inside every call.  The pool below is created once per handle,
// its workers sleep on a condition variable between calls, and — when the device's PCI function exposes
// local_cpulist in sysfs — they are pinned to the CPUs of the GPU's NUMA node, so the packed column and the page-locked
// staging they first touch sit next to the link they cross.  Moves bytes; computes nothing.
#pragma once

#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace tkb {

// CPUs local to a PCI device ("0000:3b:00.0"), from /sys/bus/pci/devices/<bdf>/local_cpulist ("0-15,32-47").
inline bool pci_local_cpus(const char* bdf, cpu_set_t* out) {
    char path[256], line[4096];
    char low[64];
    size_t n = strlen(bdf) < sizeof(low) - 1 ? strlen(bdf) : sizeof(low) - 1;
    for (size_t i = 0; i < n; ++i) low[i] = (char)((bdf[i] >= 'A' && bdf[i] <= 'F') ? bdf[i] + 32 : bdf[i]);
    low[n] = 0;
    snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/local_cpulist", low);
    FILE* f = fopen(path, "r");
    if (!f) return false;
    const bool ok = fgets(line, sizeof(line), f) != nullptr;
    fclose(f);
    if (!ok) return false;
    CPU_ZERO(out);
    int count = 0;
    for (char* p = line; *p && *p != '\n';) {
        char* end;
        long a = strtol(p, &end, 10);
        if (end == p) break;
        long b = a;
        if (*end == '-') { p = end + 1; b = strtol(p, &end, 10); if (end == p) break; }
        for (long c = a; c <= b && c < CPU_SETSIZE; ++c) { CPU_SET((int)c, out); ++count; }
        p = (*end == ',') ? end + 1 : end;
    }
    return count > 0;
}

class HostPool {
public:
    // `workers` threads beside the caller; affinity (optional): CPUs the workers are bound to
    explicit HostPool(unsigned workers, const cpu_set_t* affinity = nullptr) {
        for (unsigned i = 0; i < workers; ++i) {
            try {
                th_.emplace_back([this, i] { loop(i + 1); });
            } catch (...) {
                break;                                       // fewer workers: parallel_for still covers every part
            }
            if (affinity) pthread_setaffinity_np(th_.back().native_handle(), sizeof(cpu_set_t), affinity);
        }
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : th_) t.join();
    }
    HostPool(const HostPool&) = delete;
    HostPool& operator=(const HostPool&) = delete;

    unsigned parts() const { return (unsigned)th_.size() + 1; }

    // fn(part, parts) for part in [0, parts): part 0 on the calling thread, the others on the workers; returns when all are done
    void parallel_for(const std::function<void(unsigned, unsigned)>& fn) {
        const unsigned P = parts();
        {
            std::lock_guard<std::mutex> lk(m_);
            fn_ = &fn; pending_ = P - 1; ++epoch_;
        }
        cv_.notify_all();
        fn(0, P);
        std::unique_lock<std::mutex> lk(m_);
        done_.wait(lk, [this] { return pending_ == 0; });
        fn_ = nullptr;
    }

private:
    void loop(unsigned part) {
        unsigned long seen = 0;
        for (;;) {
            const std::function<void(unsigned, unsigned)>* fn;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [&] { return stop_ || epoch_ != seen; });
                if (stop_) return;
                seen = epoch_;
                fn = fn_;
            }
            (*fn)(part, parts());
            {
                std::lock_guard<std::mutex> lk(m_);
                if (--pending_ == 0) done_.notify_one();
            }
        }
    }
    std::vector<std::thread> th_;
    std::mutex m_;
    std::condition_variable cv_, done_;
    const std::function<void(unsigned, unsigned)>* fn_ = nullptr;
    unsigned pending_ = 0;
    unsigned long epoch_ = 0;
    bool stop_ = false;
};

}  // namespace tkb
