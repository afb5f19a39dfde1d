// shard.cu — tk_b200_shard_*: one process, several B200s, frames sharded by contiguous block.
//
// The reference is a single C process that owns ONE device per dispatcher
// (src/gpu/cuda/tk_cuda_dispatch.h:47-49 `int device_id`, tk_cuda_dispatch.cu:163-247) and feeds it from one host
// thread; it has no multi-device form.  Frames (pre-process) and head tensors (post-process) are independent units —
// no cross-frame state exists on this path — so a batch shards with NO exchange step: device g owns frames
// [g*B/G, (g+1)*B/G), processes them through its own tk_b200_pipeline on its own streams, and writes detections,
// counts and tensors straight into the caller's host arrays at the frames' own positions.  The "gather" is therefore
// already done when the call returns; tk_b200_shard_gather only compacts the padded per-frame lists by a prefix sum
// of the counts.  No NCCL, nothing crosses NVLink.
//
// One persistent host thread per device (created with the handle, parked on a condition variable between calls,
// bound to the CPUs of its GPU's NUMA node when sysfs names them) drives that device's pipeline; the caller's thread
// only hands out the blocks and waits.
#include "tk_common.cuh"

#include <condition_variable>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "host_pool.h"

namespace tkb {

struct ShardJob {
    const void* h_frames;
    const float* h_logits;
    uint32_t n;
    void* h_tensor;
    tk_b200_candidate_t* h_dets;
    uint32_t* h_ndets;
};

struct ShardWorker {
    int device = -1;
    tk_b200_pipeline_t* pipe = nullptr;
    std::thread th;
    ShardJob job{};
    tk_error_code_t rc = TK_SUCCESS;
    char detail[256] = {0};
    uint32_t over_frames = 0, over_max = 0;
};

}  // namespace tkb

struct tk_b200_shard_s {
    std::vector<tkb::ShardWorker> w;
    tk_b200_pipeline_config_t cfg;
    size_t frame_bytes = 0, logit_bytes = 0, tensor_bytes = 0;
    uint32_t per_device_batch = 0;
    std::mutex m;
    std::condition_variable cv, done;
    unsigned long epoch = 0;
    unsigned pending = 0;
    bool stop = false;
    std::mutex call_mu;                 // one process_host at a time per handle (as the reference's pipeline, tk_vision_pipeline.h:308-311)
};

using namespace tkb;

static void shard_loop(tk_b200_shard_s* s, unsigned idx) {
    ShardWorker& me = s->w[idx];
    unsigned long seen = 0;
    for (;;) {
        {
            std::unique_lock<std::mutex> lk(s->m);
            s->cv.wait(lk, [&] { return s->stop || s->epoch != seen; });
            if (s->stop) return;
            seen = s->epoch;
        }
        me.rc = TK_SUCCESS; me.detail[0] = 0; me.over_frames = me.over_max = 0;
        if (me.job.n) {
            me.rc = tk_b200_pipeline_process_host(me.pipe, me.job.h_frames, me.job.h_logits, me.job.n, me.job.h_tensor,
                                                  me.job.h_dets, me.job.h_ndets);
            if (me.rc != TK_SUCCESS) snprintf(me.detail, sizeof(me.detail), "%s", tk_error_get_detail());
            else (void)tk_b200_pipeline_candidate_overflow(me.pipe, &me.over_frames, &me.over_max);
        }
        {
            std::lock_guard<std::mutex> lk(s->m);
            if (--s->pending == 0) s->done.notify_one();
        }
    }
}

extern "C" void tk_b200_shard_destroy(tk_b200_shard_t** ps) {
    if (!ps || !*ps) return;
    tk_b200_shard_s* s = *ps;
    {
        std::lock_guard<std::mutex> lk(s->m);
        s->stop = true;
    }
    s->cv.notify_all();
    for (auto& w : s->w) if (w.th.joinable()) w.th.join();
    for (auto& w : s->w) if (w.pipe) tk_b200_pipeline_destroy(&w.pipe);
    delete s;
    *ps = nullptr;
}

// device_ids[num_devices]: the devices to shard over (an id may repeat: two pipelines on one device, which is how the
// harness is exercised on a single-GPU box).  config: as tk_b200_pipeline_create; device_id is ignored and max_batch is
// the WHOLE job's batch — each device's pipeline is sized for its block, ceil(max_batch / num_devices).
extern "C" tk_error_code_t tk_b200_shard_create(tk_b200_shard_t** out, const int* device_ids, uint32_t num_devices,
                                                const tk_b200_pipeline_config_t* config) {
    if (!out || !device_ids || !config || num_devices == 0 || num_devices > 64 || config->max_batch == 0) return TK_ERROR_INVALID_ARGUMENT;
    *out = nullptr;
    tk_b200_shard_s* s = new (std::nothrow) tk_b200_shard_s();
    if (!s) return TK_ERROR_OUT_OF_MEMORY;
    s->cfg = *config;
    s->per_device_batch = (config->max_batch + num_devices - 1) / num_devices;
    try {
        s->w.resize(num_devices);
    } catch (...) {
        delete s;
        return TK_ERROR_OUT_OF_MEMORY;
    }
    tk_error_code_t rc = TK_SUCCESS;
    for (uint32_t g = 0; g < num_devices && rc == TK_SUCCESS; ++g) {
        tk_b200_pipeline_config_t c = *config;
        c.device_id = device_ids[g];
        c.max_batch = s->per_device_batch;
        s->w[g].device = device_ids[g];
        rc = tk_b200_pipeline_create(&s->w[g].pipe, &c);
    }
    if (rc == TK_SUCCESS) {
        s->frame_bytes = tk_b200_pipeline_frame_bytes(s->w[0].pipe);
        s->tensor_bytes = tk_b200_pipeline_tensor_bytes(s->w[0].pipe);
        s->logit_bytes = (size_t)config->num_proposals * (5 + (size_t)config->num_classes) * sizeof(float);
        for (uint32_t g = 0; g < num_devices; ++g) {
            try {
                s->w[g].th = std::thread(shard_loop, s, g);
            } catch (...) {
                rc = TK_ERROR_OUT_OF_MEMORY;
                break;
            }
            cpu_set_t cpus;
            char bdf[32] = {0};
            if (cudaDeviceGetPCIBusId(bdf, (int)sizeof(bdf), device_ids[g]) == cudaSuccess && pci_local_cpus(bdf, &cpus))
                pthread_setaffinity_np(s->w[g].th.native_handle(), sizeof(cpu_set_t), &cpus);
            else
                cudaGetLastError();
        }
    }
    if (rc != TK_SUCCESS) { tk_b200_shard_destroy(&s); return rc; }
    *out = s;
    return TK_SUCCESS;
}

extern "C" uint32_t tk_b200_shard_num_devices(const tk_b200_shard_t* s) { return s ? (uint32_t)s->w.size() : 0u; }

// Block of device g: [g*n/G, (g+1)*n/G) in integer arithmetic — the blocks tile [0, n) and differ by at most one frame.
extern "C" void tk_b200_shard_block(uint32_t num_frames, uint32_t device_index, uint32_t num_devices, uint32_t* out_first, uint32_t* out_count) {
    const unsigned long long n = num_frames, g = device_index, G = num_devices ? num_devices : 1;
    const uint32_t lo = (uint32_t)(n * g / G), hi = (uint32_t)(n * (g + 1) / G);
    if (out_first) *out_first = lo;
    if (out_count) *out_count = hi - lo;
}

// src/vision/tk_object_detector.c:212-219 on every device's pipeline.
extern "C" tk_error_code_t tk_b200_shard_update_thresholds(tk_b200_shard_t* s, float conf, float iou) {
    if (!s) return TK_ERROR_INVALID_ARGUMENT;
    for (auto& w : s->w) {
        const tk_error_code_t rc = tk_b200_pipeline_update_thresholds(w.pipe, conf, iou);
        if (rc != TK_SUCCESS) return rc;
    }
    return TK_SUCCESS;
}

// HOST buffers of the whole job, laid out as for tk_b200_pipeline_process_host ([n] frames, [n] head tensors,
// [n][max_detections] detections, [n] counts, optional [n] tensors).  Every device processes its block at once; the
// call returns when all of them have (first error wins; the other devices still finish their blocks, so no copy is
// left in flight on the caller's buffers).
extern "C" tk_error_code_t tk_b200_shard_process_host(tk_b200_shard_t* s, const void* h_frames, const float* h_logits, uint32_t n,
                                                      void* h_tensor, tk_b200_candidate_t* h_dets, uint32_t* h_ndets) {
    if (!s || (!h_frames && !h_logits)) return TK_ERROR_INVALID_ARGUMENT;
    if (h_logits && (!h_dets || !h_ndets)) return TK_ERROR_INVALID_ARGUMENT;
    if (n == 0) return TK_SUCCESS;
    if (n > s->cfg.max_batch) return TK_ERROR_BUFFER_TOO_SMALL;
    std::lock_guard<std::mutex> call(s->call_mu);
    const uint32_t G = (uint32_t)s->w.size();
    const uint32_t max_det = s->cfg.max_detections ? s->cfg.max_detections : 500u;
    for (uint32_t g = 0; g < G; ++g) {
        uint32_t lo, cnt;
        tk_b200_shard_block(n, g, G, &lo, &cnt);
        ShardJob& j = s->w[g].job;
        j.n = cnt;
        j.h_frames = h_frames ? static_cast<const uint8_t*>(h_frames) + (size_t)lo * s->frame_bytes : nullptr;
        j.h_logits = h_logits ? reinterpret_cast<const float*>(reinterpret_cast<const uint8_t*>(h_logits) + (size_t)lo * s->logit_bytes) : nullptr;
        j.h_tensor = h_tensor ? static_cast<uint8_t*>(h_tensor) + (size_t)lo * s->tensor_bytes : nullptr;
        j.h_dets = h_dets ? h_dets + (size_t)lo * max_det : nullptr;
        j.h_ndets = h_ndets ? h_ndets + lo : nullptr;
    }
    {
        std::lock_guard<std::mutex> lk(s->m);
        s->pending = G;
        ++s->epoch;
    }
    s->cv.notify_all();
    {
        std::unique_lock<std::mutex> lk(s->m);
        s->done.wait(lk, [&] { return s->pending == 0; });
    }
    for (uint32_t g = 0; g < G; ++g) {
        if (s->w[g].rc != TK_SUCCESS) {
            tk_error_set_detail("device %d (block %u of %u): %s", s->w[g].device, g, G, s->w[g].detail);
            return s->w[g].rc;
        }
    }
    return TK_SUCCESS;
}

// As tk_b200_pipeline_candidate_overflow, summed over the devices of the last call.
extern "C" tk_error_code_t tk_b200_shard_candidate_overflow(tk_b200_shard_t* s, uint32_t* out_frames, uint32_t* out_max_count) {
    if (!s) return TK_ERROR_INVALID_ARGUMENT;
    uint32_t f = 0, m = 0;
    for (auto& w : s->w) { f += w.over_frames; if (w.over_max > m) m = w.over_max; }
    if (out_frames) *out_frames = f;
    if (out_max_count) *out_max_count = m;
    return TK_SUCCESS;
}

// The final host gather ("prefix-sum of counts, memcpy"): padded per-frame lists -> one flat list in global
// frame order; out_offsets[n + 1], frame f's detections are out_flat[out_offsets[f] .. out_offsets[f + 1]).  out_flat
// may be NULL to size it (returns the total through out_offsets[n]).  Moves records; computes nothing.
extern "C" tk_error_code_t tk_b200_shard_gather(const tk_b200_candidate_t* h_dets, const uint32_t* h_counts, uint32_t n,
                                                uint32_t max_detections, tk_b200_candidate_t* out_flat, uint64_t* out_offsets) {
    if (!h_counts || !out_offsets || (n && !h_dets && out_flat)) return TK_ERROR_INVALID_ARGUMENT;
    uint64_t run = 0;
    for (uint32_t f = 0; f < n; ++f) {
        out_offsets[f] = run;
        const uint32_t k = h_counts[f] < max_detections ? h_counts[f] : max_detections;
        if (out_flat && k) memcpy(out_flat + run, h_dets + (size_t)f * max_detections, (size_t)k * sizeof(tk_b200_candidate_t));
        run += k;
    }
    out_offsets[n] = run;
    return TK_SUCCESS;
}
