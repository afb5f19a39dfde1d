// pipeline.cu — batched pre-process + decode + class-wise NMS behind one handle.
//
// The reference runs this path one frame per call inside
// tk_object_detector_detect (src/vision/tk_object_detector.c:182-203:
// preprocess_frame :223-259 with a synchronous pageable cudaMemcpy :229,
// postprocess_detections :303-368).  Here a batch of frames of one geometry goes
// through three launches (preprocess_kernel, decode_kernel, nms_kernel), and the
// host-buffer form cuts the batch into chunks so that the upload of chunk k+1,
// the kernels of chunk k and the download of chunk k-1 overlap on three streams
// ordered by events (the reference's dispatcher creates three streams and never
// orders them, src/gpu/cuda/tk_cuda_dispatch.cu:107-137).
//
// The host side moves bytes (chunked copies, packing the objectness column of page-locked logits into a dense array);
// it computes nothing: thresholds, boxes, scores and suppression are all decided on the device.
#include "tk_common.cuh"

#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <new>

#include "host_pool.h"

namespace tkb {

static constexpr int kSlots = 3;            // in-flight chunks of process_host

struct DevScope {
    int prev = -1; bool ok = false;
    explicit DevScope(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        ok = cudaSetDevice(dev) == cudaSuccess;
        if (!ok) cudaGetLastError();
    }
    ~DevScope() { if (prev >= 0 ) cudaSetDevice(prev); }
};

}  // namespace tkb

struct tk_b200_pipeline_s {
    tk_b200_pipeline_config_t cfg;
    size_t frame_bytes, logit_bytes, tensor_bytes;     // per frame
    uint32_t chunk;
    cudaStream_t s_up, s_compute, s_down, s_post;
    cudaStream_t s_side;                               // process_device: the fused post-process kernel runs here, under the pre-process kernel
    cudaEvent_t ev_post, ev_fork, ev_join;
    int fused;                                         // fuse_postprocess: 0 / 1 separate kernels (default), 2 co-resident, 3 cluster-fused
    void* d_post_ws; size_t post_ws_bytes;             // its global NMS work area (frames above 512 candidates)
    uint32_t* d_scan;                                  // [max_batch][ceil(P / 32)] ballot words of decode's pass 1, written by the pre-process kernel's scanner warps
    int scan;                                          // 1: process_device runs that pass inside the pre-process kernel
    cudaEvent_t ev_uploaded[tkb::kSlots], ev_computed[tkb::kSlots], ev_downloaded[tkb::kSlots];
    uint8_t* d_frames[tkb::kSlots];                    // [chunk][frame_bytes (+ slack)]
    float*   d_logits[tkb::kSlots];                    // [chunk][P][5+C]
    uint8_t* d_tensor;                                 // [max_batch][tensor_bytes]
    tk_b200_candidate_t* d_cands;                      // [max_batch][max_candidates]
    uint32_t* d_counts;                                // [max_batch]
    tk_b200_candidate_t* d_dets;                       // [max_batch][max_detections]
    uint32_t* d_ndets;                                 // [max_batch]
    float* h_obj;                                      // [max_batch][P] page-locked: objectness column packed on the host
    float* d_obj;                                      // [max_batch][P]
    tkb::HostPool* pool;                               // host threads that pack it (created on first use, NUMA-local to the GPU)
    uint32_t* h_cand_counts;                           // [max_batch] page-locked: rows over the threshold per frame of the last host call
    uint32_t counts_valid, over_frames, over_max;      // ... and what they say about the candidate cap
    std::mutex thr_mu;                                 // guards thr_conf / thr_iou (update_thresholds vs a running call)
    float thr_conf, thr_iou;
    // Optional per-stage timing of process_device: 4 events per call (before pre,
    // after pre, after decode, after nms) on the caller's stream.
    int timing_on;
    uint32_t timing_calls, timing_cap;
    cudaEvent_t* timing_ev;                            // [timing_cap][4]
    uint8_t* timing_mode;                              // [timing_cap] 0 three kernels in a row, 1 fused post-process behind, 2 beside the pre-process kernel
};

using namespace tkb;

static size_t frame_bytes_of(const tk_b200_pipeline_config_t& c) {
    uint32_t bpp = (c.input_format == TK_PIXEL_FORMAT_RGB8 || c.input_format == TK_PIXEL_FORMAT_BGR8) ? 3
                   : (c.input_format == TK_PIXEL_FORMAT_RGBA8 ? 4 : 1);
    size_t stride = c.input_stride_bytes ? c.input_stride_bytes : (size_t)c.input_width * bpp;
    size_t b = stride * c.input_height;
    if (c.input_format == TK_PIXEL_FORMAT_I420) b += 2 * (stride / 2) * (size_t)(c.input_height / 2);
    return b;
}

extern "C" void tk_b200_pipeline_destroy(tk_b200_pipeline_t** pp) {
    if (!pp || !*pp) return;
    tk_b200_pipeline_t* p = *pp;
    {
        DevScope scope(p->cfg.device_id);
        if (p->s_up) { cudaStreamSynchronize(p->s_up); }
        if (p->s_compute) { cudaStreamSynchronize(p->s_compute); }
        if (p->s_down) { cudaStreamSynchronize(p->s_down); }
        if (p->s_post) { cudaStreamSynchronize(p->s_post); }
        if (p->s_side) { cudaStreamSynchronize(p->s_side); }
        for (int i = 0; i < kSlots; ++i) {
            if (p->d_frames[i]) cudaFree(p->d_frames[i]);
            if (p->d_logits[i]) cudaFree(p->d_logits[i]);
            if (p->ev_uploaded[i]) cudaEventDestroy(p->ev_uploaded[i]);
            if (p->ev_computed[i]) cudaEventDestroy(p->ev_computed[i]);
            if (p->ev_downloaded[i]) cudaEventDestroy(p->ev_downloaded[i]);
        }
        if (p->d_tensor) cudaFree(p->d_tensor);
        if (p->d_cands) cudaFree(p->d_cands);
        if (p->d_counts) cudaFree(p->d_counts);
        if (p->d_dets) cudaFree(p->d_dets);
        if (p->d_ndets) cudaFree(p->d_ndets);
        if (p->d_obj) cudaFree(p->d_obj);
        if (p->h_obj) cudaFreeHost(p->h_obj);
        if (p->h_cand_counts) cudaFreeHost(p->h_cand_counts);
        delete p->pool;
        for (uint32_t i = 0; p->timing_ev && i < p->timing_cap * 4; ++i) if (p->timing_ev[i]) cudaEventDestroy(p->timing_ev[i]);
        free(p->timing_ev); free(p->timing_mode);
        if (p->s_up) cudaStreamDestroy(p->s_up);
        if (p->s_compute) cudaStreamDestroy(p->s_compute);
        if (p->s_down) cudaStreamDestroy(p->s_down);
        if (p->s_post) cudaStreamDestroy(p->s_post);
        if (p->s_side) cudaStreamDestroy(p->s_side);
        if (p->ev_post) cudaEventDestroy(p->ev_post);
        if (p->ev_fork) cudaEventDestroy(p->ev_fork);
        if (p->ev_join) cudaEventDestroy(p->ev_join);
        if (p->d_post_ws) cudaFree(p->d_post_ws);
        if (p->d_scan) cudaFree(p->d_scan);
        cudaGetLastError();
    }
    delete p;
    *pp = nullptr;
}

extern "C" tk_error_code_t tk_b200_pipeline_create(tk_b200_pipeline_t** out, const tk_b200_pipeline_config_t* cfg) {
    if (!out || !cfg) return TK_ERROR_INVALID_ARGUMENT;
    *out = nullptr;
    if (cfg->max_batch == 0) return TK_ERROR_INVALID_ARGUMENT;
    if (cfg->input_width < 2 || cfg->input_height < 2 || cfg->tensor_width == 0 || cfg->tensor_height == 0)
        return TK_ERROR_INVALID_DIMENSIONS;
    if (cfg->input_format != TK_PIXEL_FORMAT_RGB8 && cfg->input_format != TK_PIXEL_FORMAT_RGBA8 &&
        cfg->input_format != TK_PIXEL_FORMAT_I420 && cfg->input_format != TK_PIXEL_FORMAT_BGR8)
        return TK_ERROR_INVALID_ARGUMENT;
    if (cfg->max_candidates > 4096) return TK_ERROR_INVALID_ARGUMENT;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return TK_ERROR_GPU_DEVICE_ENUMERATION_FAILED; }
    if (count == 0 || cfg->device_id < 0 || cfg->device_id >= count) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    DevScope scope(cfg->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    if (!device_info()) return TK_ERROR_GPU_DEVICE_UNSUPPORTED;

    tk_b200_pipeline_t* p = new (std::nothrow) tk_b200_pipeline_t();
    if (!p) return TK_ERROR_OUT_OF_MEMORY;
    p->cfg = *cfg;
    p->thr_conf = cfg->confidence_threshold; p->thr_iou = cfg->iou_threshold;
    if (p->cfg.max_candidates == 0) p->cfg.max_candidates = 4096;
    if (p->cfg.max_detections == 0) p->cfg.max_detections = 500;      // TK_OBJECT_DETECTOR_MAX_DETECTIONS, tk_object_detector.c:39
    p->chunk = cfg->chunk_frames ? cfg->chunk_frames : 8;
    if (p->chunk > cfg->max_batch) p->chunk = cfg->max_batch;
    p->frame_bytes = frame_bytes_of(*cfg);
    p->logit_bytes = (size_t)cfg->num_proposals * (5 + (size_t)cfg->num_classes) * sizeof(float);
    p->tensor_bytes = (size_t)cfg->tensor_width * cfg->tensor_height * 3 * (cfg->tensor_type == TK_B200_OUT_F16 ? 2 : 4);

    bool ok = cudaStreamCreateWithFlags(&p->s_up, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->s_compute, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->s_down, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&p->s_post, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&p->ev_post, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming) == cudaSuccess;
    if (ok) {
        // the side stream gets the highest priority: its few small CTAs should be placed before the pre-process grid fills the SMs
        int lo = 0, hi = 0;
        if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess) { cudaGetLastError(); hi = 0; }
        ok = cudaStreamCreateWithPriority(&p->s_side, cudaStreamNonBlocking, hi) == cudaSuccess;
    }
    for (int i = 0; ok && i < kSlots; ++i) {
        ok = cudaEventCreateWithFlags(&p->ev_uploaded[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&p->ev_computed[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&p->ev_downloaded[i], cudaEventDisableTiming) == cudaSuccess;
    }
    if (!ok) { cudaGetLastError(); tk_b200_pipeline_destroy(&p); return TK_ERROR_GPU_STREAM_CREATION_FAILED; }

    const size_t mb = cfg->max_batch;
    // +64 B: the row ring's 16-byte bulk copies may run past the last row of the last frame.
    for (int i = 0; ok && i < kSlots; ++i) {
        ok = cudaMalloc(reinterpret_cast<void**>(&p->d_frames[i]), p->chunk * p->frame_bytes + 64) == cudaSuccess;
        if (ok && p->logit_bytes) ok = cudaMalloc(reinterpret_cast<void**>(&p->d_logits[i]), p->chunk * p->logit_bytes) == cudaSuccess;
    }
    ok = ok && cudaMalloc(reinterpret_cast<void**>(&p->d_tensor), mb * p->tensor_bytes) == cudaSuccess;
    ok = ok && cudaMalloc(reinterpret_cast<void**>(&p->d_cands), mb * p->cfg.max_candidates * sizeof(tk_b200_candidate_t)) == cudaSuccess;
    ok = ok && cudaMalloc(reinterpret_cast<void**>(&p->d_counts), mb * sizeof(uint32_t)) == cudaSuccess;
    ok = ok && cudaMalloc(reinterpret_cast<void**>(&p->d_dets), mb * p->cfg.max_detections * sizeof(tk_b200_candidate_t)) == cudaSuccess;
    ok = ok && cudaMalloc(reinterpret_cast<void**>(&p->d_ndets), mb * sizeof(uint32_t)) == cudaSuccess;
    ok = ok && cudaHostAlloc(reinterpret_cast<void**>(&p->h_cand_counts), mb * sizeof(uint32_t), cudaHostAllocDefault) == cudaSuccess;
    // decode's pass 1 inside the pre-process kernel (preprocess.cu, ScanK): OFF unless TK_B200_SCAN=1.  Measured on
    // 64 x 1080p (profiles/r2_experiments.md): decode_kernel 41.9 -> 18.2 us, but the scattered 64-byte fetches of the
    // scan cost the pre-process kernel's row stream 27.5 us (119.2 -> 146.7 us) — the scan is DRAM work, not idle
    // latency, and takes its ~23 us of the memory system wherever it runs; the step got 3.6 us slower.
    p->scan = 0;
    if (const char* e = getenv("TK_B200_SCAN")) p->scan = atoi(e) != 0;
    if (ok && p->scan && cfg->num_proposals)
        ok = cudaMalloc(reinterpret_cast<void**>(&p->d_scan), mb * ((cfg->num_proposals + 31) / 32) * sizeof(uint32_t)) == cudaSuccess;
    // decode + NMS as one kernel (post_mode); TK_B200_POST_FUSED overrides fuse_postprocess (experiments).
    p->fused = cfg->fuse_postprocess;
    if (const char* e = getenv("TK_B200_POST_FUSED")) p->fused = atoi(e);
    if (ok && (p->fused == 2 || p->fused == 3)) {
        p->post_ws_bytes = tk_b200_postprocess_fused_workspace_bytes(p->cfg.max_candidates, cfg->max_batch, p->fused == 2 ? 1 : 0);
        if (p->post_ws_bytes) ok = cudaMalloc(&p->d_post_ws, p->post_ws_bytes) == cudaSuccess;
    }
    if (!ok) {
        tk_error_set_detail("pipeline buffers: %s", cudaGetErrorString(cudaGetLastError()));
        tk_b200_pipeline_destroy(&p);
        return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
    }
    *out = p;
    return TK_SUCCESS;
}

// src/vision/tk_object_detector.c:212-219; the reference's pipeline calls it under its mutex
// (src/vision/tk_vision_pipeline.c:131-150).  Safe against a process_* call running on another thread: the new values
// are picked up by the next call.
extern "C" tk_error_code_t tk_b200_pipeline_update_thresholds(tk_b200_pipeline_t* p, float conf, float iou) {
    if (!p) return TK_ERROR_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lk(p->thr_mu);
    p->thr_conf = conf;
    p->thr_iou = iou;
    return TK_SUCCESS;
}

// After tk_b200_pipeline_process_host[_frames]: frames of that call whose rows over the confidence threshold
// outnumbered max_candidates (their candidate lists were cut to the first max_candidates rows by proposal index
// before NMS — the reference keeps every candidate, src/vision/tk_object_detector.c:339-355), and the largest count.
extern "C" tk_error_code_t tk_b200_pipeline_candidate_overflow(tk_b200_pipeline_t* p, uint32_t* out_frames, uint32_t* out_max_count) {
    if (!p) return TK_ERROR_INVALID_ARGUMENT;
    if (out_frames) *out_frames = p->over_frames;
    if (out_max_count) *out_max_count = p->over_max;
    return TK_SUCCESS;
}
// process_device leaves the per-frame counts of rows over the threshold on the device: [n] u32, valid once the call's
// work on the stream has completed.
extern "C" const uint32_t* tk_b200_pipeline_candidate_counts_device_ptr(tk_b200_pipeline_t* p) { return p ? p->d_counts : nullptr; }

// Host threads + page-locked column for the objectness packing of process_host, created on first use.  The workers
// are bound to the CPUs of the GPU's NUMA node when sysfs names them, and the column is allocated (first touched) by
// one of them, so the packed bytes sit on the memory next to the link they cross.
static tk_error_code_t ensure_pack_resources(tk_b200_pipeline_t* p) {
    if (!p->pool) {
        const unsigned hw = std::thread::hardware_concurrency();
        unsigned T = hw >= 16 ? 4u : (hw >= 4 ? 2u : 1u);
        if (const char* e = getenv("TK_B200_PACK_THREADS")) { const int v = atoi(e); if (v >= 1 && v <= 64) T = (unsigned)v; }
        cpu_set_t cpus;
        char bdf[32] = {0};
        const bool local = cudaDeviceGetPCIBusId(bdf, (int)sizeof(bdf), p->cfg.device_id) == cudaSuccess && tkb::pci_local_cpus(bdf, &cpus);
        if (!local) cudaGetLastError();
        p->pool = new (std::nothrow) tkb::HostPool(T - 1, local ? &cpus : nullptr);
        if (!p->pool) return TK_ERROR_OUT_OF_MEMORY;
    }
    if (!p->h_obj || !p->d_obj) {
        const size_t ob = (size_t)p->cfg.max_batch * p->cfg.num_proposals * sizeof(float);
        cudaError_t e_host = cudaSuccess;
        if (!p->h_obj) {
            const int dev = p->cfg.device_id;
            float** slot = &p->h_obj;
            p->pool->parallel_for([&](unsigned part, unsigned parts) {       // allocate from the last (bound) worker
                if (part != parts - 1) return;
                cudaSetDevice(dev);
                e_host = cudaHostAlloc(reinterpret_cast<void**>(slot), ob, cudaHostAllocDefault);
            });
        }
        if (e_host != cudaSuccess || (!p->d_obj && cudaMalloc(reinterpret_cast<void**>(&p->d_obj), ob) != cudaSuccess)) {
            tk_error_set_detail("pipeline objectness staging: %s", cudaGetErrorString(cudaGetLastError()));
            return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
        }
    }
    return TK_SUCCESS;
}

extern "C" void* tk_b200_pipeline_tensor_device_ptr(tk_b200_pipeline_t* p) { return p ? p->d_tensor : nullptr; }
extern "C" size_t tk_b200_pipeline_frame_bytes(const tk_b200_pipeline_t* p) { return p ? p->frame_bytes : 0; }
extern "C" size_t tk_b200_pipeline_tensor_bytes(const tk_b200_pipeline_t* p) { return p ? p->tensor_bytes : 0; }

static tk_b200_preprocess_batch_params_t pre_params(const tk_b200_pipeline_t* p, const void* d_frames, uint32_t n, void* d_tensor) {
    const tk_b200_pipeline_config_t& c = p->cfg;
    tk_b200_preprocess_batch_params_t b{};
    b.d_input = d_frames; b.input_frame_pitch_bytes = p->frame_bytes; b.num_frames = n;
    b.input_width = c.input_width; b.input_height = c.input_height; b.input_stride_bytes = c.input_stride_bytes;
    b.input_format = c.input_format;
    b.d_output = d_tensor; b.output_width = c.tensor_width; b.output_height = c.tensor_height;
    b.output_type = c.tensor_type; b.resize_mode = c.resize_mode; b.pad_value = c.pad_value; b.arith = c.arith;
    b.mean = c.mean; b.std_dev = c.std_dev;
    return b;
}

// How `n` frames are post-processed: 0 decode_kernel then nms_kernel (the default: fastest as measured);
// 3 ONE kernel, a cluster of 8 CTAs per frame that decodes and then runs the frame's NMS
// (tk_kernels_postprocess_fused layout 0); 2 the persistent co-resident form (layout 1) forked onto the side stream
// beside the pre-process kernel.  Both fused forms are bit-identical to the default and selectable, never automatic.
static int post_mode(const tk_b200_pipeline_t* p, uint32_t n) {
    (void)n;
    return (p->fused == 2 || p->fused == 3) ? p->fused : 0;
}

// The launches for `n` frames whose inputs are in HBM: the pre-process kernel on `st`, then the post-process on `st`
// (mode 3: one launch; mode 0: two), or — mode 2 — post_kernel on the side stream, forked off `st` before the
// pre-process launch and joined after it.
// tev (stage timing): [0] start, [1] pre-process done, [2] decode done (fused: the whole post-process done; mode 2:
// recorded on the side stream), [3] end.  *out_mode: 0 separate kernels, 1 fused in-stream, 2 fused beside.
static tk_error_code_t run_kernels(tk_b200_pipeline_t* p, const void* d_frames, const float* d_logits, uint32_t n,
                                   void* d_tensor, tk_b200_candidate_t* d_cands, uint32_t* d_counts,
                                   tk_b200_candidate_t* d_dets, uint32_t* d_ndets, cudaStream_t st,
                                   cudaEvent_t* tev = nullptr, const float* d_obj = nullptr, uint8_t* out_mode = nullptr,
                                   bool pdl_pre = false) {
    const tk_b200_pipeline_config_t& c = p->cfg;
    tk_error_code_t rc;
    if (out_mode) *out_mode = 0;
    if (d_logits && (!d_dets || !d_ndets)) return TK_ERROR_INVALID_ARGUMENT;
    if (d_frames && !d_tensor) return TK_ERROR_INVALID_ARGUMENT;
    if (d_logits && c.restore_coordinates && c.resize_mode != TK_B200_RESIZE_STRETCH)
        return TK_ERROR_NOT_IMPLEMENTED;   // a letterboxed tensor would also need the pad offsets, which the reference
                                           // never defines (convert_to_original_coordinates is empty)
    int mode = d_logits ? post_mode(p, n) : 0;
    const bool fork = mode == 2 && d_frames;                   // something to run beside
    auto fused_params = [&](int layout) {
        tk_b200_post_params_t q{};
        q.d_logits = d_logits; q.num_frames = n; q.num_proposals = c.num_proposals; q.num_classes = c.num_classes;
        q.confidence_threshold = c.confidence_threshold; q.activation = c.activation;
        q.d_candidates = d_cands; q.d_counts = d_counts; q.max_candidates = c.max_candidates;
        q.iou_threshold = c.iou_threshold; q.max_detections = c.max_detections; q.class_agnostic = c.class_agnostic;
        q.d_detections = d_dets; q.d_num_detections = d_ndets;
        if (c.restore_coordinates) {
            q.restore_src_width = c.input_width; q.restore_src_height = c.input_height;
            q.restore_in_width = c.tensor_width; q.restore_in_height = c.tensor_height;
        }
        q.d_workspace = p->d_post_ws; q.workspace_bytes = p->post_ws_bytes; q.layout = layout;
        return q;
    };
    if (tev) cudaEventRecord(tev[0], st);
    bool post_done = false;
    if (fork) {
        TK_CUDA_TRY(cudaEventRecord(p->ev_fork, st), TK_ERROR_GPU_CUDA_ERROR);
        TK_CUDA_TRY(cudaStreamWaitEvent(p->s_side, p->ev_fork, 0), TK_ERROR_GPU_CUDA_ERROR);
        const tk_b200_post_params_t q = fused_params(1);
        rc = postprocess_fused_packed(&q, d_obj, p->s_side);
        if (rc == TK_SUCCESS) {
            if (tev) cudaEventRecord(tev[2], p->s_side);
            TK_CUDA_TRY(cudaEventRecord(p->ev_join, p->s_side), TK_ERROR_GPU_CUDA_ERROR);
            post_done = true;
            if (out_mode) *out_mode = 2;
        } else if (rc != TK_ERROR_NOT_IMPLEMENTED) {
            return rc;
        } else {
            mode = 0;                                          // a head shape its 25 KB do not hold: the two kernels, in stream
        }
    }
    // decode's pass 1 rides in the pre-process kernel when both run in this call, back to back on one stream, on
    // logits in HBM (the PACKED host path has its objectness column already)
    const bool scan = p->scan && p->d_scan && d_frames && d_logits && !post_done && mode == 0 && !d_obj;
    tk_b200_decode_params_t d{};
    if (d_logits) {
        d.d_logits = d_logits; d.num_frames = n; d.num_proposals = c.num_proposals; d.num_classes = c.num_classes;
        d.confidence_threshold = c.confidence_threshold; d.activation = c.activation;
        d.d_candidates = d_cands; d.d_counts = d_counts; d.max_candidates = c.max_candidates;
    }
    // pre-process -> decode -> NMS as a programmatic-dependent-launch chain (tk_common.cuh, pdl_chain_hint): OFF unless
    // TK_B200_PDL_CHAIN=1.  Measured twice (profiles/r2_experiments.md §7, §12): with the dependents announced at the top
    // of each kernel 182.8 us per 64-frame step against 164.8 us — the decode clusters that become resident under the
    // pre-process kernel's tail only wait there, and what they hold slows the CTAs still working; announced at the END of
    // each kernel (the form in the tree) 157.7 us against 149.4 us.  Only process_device (pdl_pre), three-kernel form.
    static const bool chain_env = [] { const char* e = getenv("TK_B200_PDL_CHAIN"); return e && atoi(e) != 0; }();
    struct HintGuard { bool& h; bool old; HintGuard(bool& r, bool v) : h(r), old(r) { h = v; } ~HintGuard() { h = old; } }
        hint(pdl_chain_hint(), chain_env && pdl_pre && d_frames && d_logits && !post_done && mode == 0);
    if (d_frames) {
        const tk_b200_preprocess_batch_params_t b = pre_params(p, d_frames, n, d_tensor);
        if ((rc = preprocess_batch_impl(&b, st, pdl_pre, scan ? &d : nullptr, scan ? p->d_scan : nullptr)) != TK_SUCCESS) return rc;
    }
    if (tev) cudaEventRecord(tev[1], st);
    if (post_done) TK_CUDA_TRY(cudaStreamWaitEvent(st, p->ev_join, 0), TK_ERROR_GPU_CUDA_ERROR);
    if (d_logits && !post_done && mode != 0) {
        const tk_b200_post_params_t q = fused_params(mode == 2 ? 1 : 0);
        rc = postprocess_fused_packed(&q, d_obj, st);
        if (rc == TK_SUCCESS) {
            post_done = true;
            if (tev) cudaEventRecord(tev[2], st);
            if (out_mode) *out_mode = 1;
        } else if (rc != TK_ERROR_NOT_IMPLEMENTED) {
            return rc;
        }
    }
    if (d_logits && !post_done) {
        if ((rc = scan ? decode_detections_bitmap(&d, p->d_scan, st) : decode_detections_packed(&d, d_obj, st)) != TK_SUCCESS) return rc;
        if (tev) cudaEventRecord(tev[2], st);
        tk_b200_nms_params_t m{};
        m.d_candidates = d_cands; m.d_counts = d_counts; m.num_frames = n; m.max_candidates = c.max_candidates;
        m.iou_threshold = c.iou_threshold; m.max_detections = c.max_detections; m.class_agnostic = c.class_agnostic;
        m.d_detections = d_dets; m.d_num_detections = d_ndets;
        if (c.restore_coordinates) {
            m.restore_src_width = c.input_width; m.restore_src_height = c.input_height;
            m.restore_in_width = c.tensor_width; m.restore_in_height = c.tensor_height;
        }
        if ((rc = tk_kernels_nms_classwise(&m, st)) != TK_SUCCESS) return rc;
    } else if (tev && !d_logits) {
        cudaEventRecord(tev[2], st);
    }
    if (tev) cudaEventRecord(tev[3], st);
    return TK_SUCCESS;
}

extern "C" tk_error_code_t tk_b200_pipeline_process_device(tk_b200_pipeline_t* p, const void* d_frames, const float* d_logits,
                                                           uint32_t n, void* d_tensor, tk_b200_candidate_t* d_dets,
                                                           uint32_t* d_ndets, cudaStream_t stream) {
    if (!p || (!d_frames && !d_logits)) return TK_ERROR_INVALID_ARGUMENT;
    if (n == 0) return TK_SUCCESS;
    if (n > p->cfg.max_batch) return TK_ERROR_BUFFER_TOO_SMALL;
    DevScope scope(p->cfg.device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    {
        std::lock_guard<std::mutex> lk(p->thr_mu);
        p->cfg.confidence_threshold = p->thr_conf;
        p->cfg.iou_threshold = p->thr_iou;
    }
    cudaEvent_t* tev = nullptr;
    uint8_t* mode = nullptr;
    if (p->timing_on && p->timing_calls < p->timing_cap) {
        mode = p->timing_mode + p->timing_calls;
        tev = p->timing_ev + 4 * (size_t)p->timing_calls++;
    }
    return run_kernels(p, d_frames, d_logits, n, d_tensor, p->d_cands, p->d_counts, d_dets, d_ndets, stream, tev, nullptr, mode,
                       /*pdl_pre=*/true);
}

// Stage timing of process_device with CUDA events on the caller's stream.
// enable(max_calls > 0) arms it for the next max_calls calls; enable(0) disarms.
extern "C" tk_error_code_t tk_b200_pipeline_stage_timing(tk_b200_pipeline_t* p, uint32_t max_calls) {
    if (!p) return TK_ERROR_INVALID_ARGUMENT;
    DevScope scope(p->cfg.device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    p->timing_on = 0; p->timing_calls = 0;
    if (max_calls == 0) return TK_SUCCESS;
    if (max_calls > p->timing_cap) {
        cudaEvent_t* ev = static_cast<cudaEvent_t*>(calloc((size_t)max_calls * 4, sizeof(cudaEvent_t)));
        uint8_t* md = static_cast<uint8_t*>(calloc(max_calls, 1));
        if (!ev || !md) { free(ev); free(md); return TK_ERROR_OUT_OF_MEMORY; }
        free(p->timing_mode); p->timing_mode = md;
        for (uint32_t i = 0; i < max_calls * 4; ++i) {
            if (i < p->timing_cap * 4) ev[i] = p->timing_ev[i];
            else if (cudaEventCreate(&ev[i]) != cudaSuccess) {
                // keep what exists: the old events stay owned by the old table
                for (uint32_t j = p->timing_cap * 4; j < i; ++j) cudaEventDestroy(ev[j]);
                cudaGetLastError(); free(ev); return TK_ERROR_GPU_CUDA_ERROR;
            }
        }
        free(p->timing_ev);
        p->timing_ev = ev; p->timing_cap = max_calls;
    }
    p->timing_on = 1;
    return TK_SUCCESS;
}

// Sums of the recorded stage durations (ms): [0] pre-process, [1] decode, [2] NMS; *out_calls = calls recorded.
// Calls that ran a fused post-process kernel: [1] = decode + NMS together and [2] = 0 (the co-resident form, mode 2,
// is timed from the fork: it runs beside the pre-process kernel, so [0] + [1] exceeds the step).  Waits for the events.
extern "C" tk_error_code_t tk_b200_pipeline_stage_times(tk_b200_pipeline_t* p, double ms_sum[3], uint32_t* out_calls) {
    if (!p || !ms_sum || !out_calls) return TK_ERROR_INVALID_ARGUMENT;
    DevScope scope(p->cfg.device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    ms_sum[0] = ms_sum[1] = ms_sum[2] = 0.0;
    *out_calls = p->timing_calls;
    for (uint32_t c = 0; c < p->timing_calls; ++c) {
        cudaEvent_t* e = p->timing_ev + 4 * (size_t)c;
        TK_CUDA_TRY(cudaEventSynchronize(e[3]), TK_ERROR_GPU_SYNCHRONIZATION_FAILED);
        if (p->timing_mode[c] != 0) {
            float a = 0.0f, b = 0.0f;
            TK_CUDA_TRY(cudaEventSynchronize(e[2]), TK_ERROR_GPU_SYNCHRONIZATION_FAILED);
            TK_CUDA_TRY(cudaEventElapsedTime(&a, e[0], e[1]), TK_ERROR_GPU_CUDA_ERROR);
            // fused in stream: behind the pre-process kernel; fused beside it: from the fork
            TK_CUDA_TRY(cudaEventElapsedTime(&b, p->timing_mode[c] == 2 ? e[0] : e[1], e[2]), TK_ERROR_GPU_CUDA_ERROR);
            ms_sum[0] += (double)a; ms_sum[1] += (double)b;
            continue;
        }
        for (int s = 0; s < 3; ++s) {
            float ms = 0.0f;
            TK_CUDA_TRY(cudaEventElapsedTime(&ms, e[s], e[s + 1]), TK_ERROR_GPU_CUDA_ERROR);
            ms_sum[s] += (double)ms;
        }
    }
    return TK_SUCCESS;
}

// Host-buffer flow.
//
// Frames: chunked three-stream flow.  Slot s = chunk % kSlots owns one frame buffer
// (and one logit buffer) and its events:
//   upload(k)   waits computed(k - kSlots)   (the slot's previous kernels are done reading)
//   compute(k)  waits uploaded(k)
//   download(k) waits computed(k)
// Outputs (tensor, detections, counts) have per-frame homes for the whole batch,
// so downloads never block later compute.
//
// Logits: decode needs 4 bytes of every 340-byte row (the objectness word) and the
// full row only for the few hundred survivors, so copying the tensor to HBM first
// moves ~50x more bytes over PCIe than the kernel needs.  When the caller's logits
// are in page-locked host memory (tk_b200_host_alloc / cudaHostAlloc: device-
// addressable under UVA) they stay there: the objectness column is packed into a
// dense page-locked array by the handle's host threads while the frame copies run,
// uploaded as one copy, and the PACKED decode reads only the surviving rows in place over
// PCIe (stage_host_logits = 2 or a batch of a frame or two: the objectness words are
// read in place as well, one PCIe request per row).  One launch for the whole batch
// on its own stream, concurrent with the frame uploads.  Pageable logits (or
// stage_host_logits = 1) are staged through HBM chunk by chunk.
//
// Frames of a sparse geometry are not copied either (see frames_in_place below).
//
// Errors: nothing returns out of the middle — a failed enqueue sets rc, stops further enqueues and falls through to
// the drain at the end, so no copy is left in flight on the caller's buffers when the call returns.
static tk_error_code_t process_host_impl(tk_b200_pipeline_t* p, const void* h_frames, const float* h_logits,
                                         const float* d_logits_all, uint32_t n, void* h_tensor,
                                         tk_b200_candidate_t* h_dets, uint32_t* h_ndets) {
    if (!p) return TK_ERROR_INVALID_ARGUMENT;
    const bool have_logits = h_logits || d_logits_all;
    if (!h_frames && !have_logits) return TK_ERROR_INVALID_ARGUMENT;
    if (have_logits && (!h_dets || !h_ndets)) return TK_ERROR_INVALID_ARGUMENT;
    if (have_logits && p->logit_bytes == 0) return TK_ERROR_INVALID_DIMENSIONS;
    if (n == 0) return TK_SUCCESS;
    if (n > p->cfg.max_batch) return TK_ERROR_BUFFER_TOO_SMALL;
    DevScope scope(p->cfg.device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    {   // thresholds as of now (tk_b200_pipeline_update_thresholds may run on another thread, as the reference's
        // pipeline does under its mutex, src/vision/tk_vision_pipeline.c:131-150)
        std::lock_guard<std::mutex> lk(p->thr_mu);
        p->cfg.confidence_threshold = p->thr_conf;
        p->cfg.iou_threshold = p->thr_iou;
    }
    const size_t det_stride = p->cfg.max_detections;
    tk_error_code_t rc = TK_SUCCESS;
    p->over_frames = 0; p->over_max = 0; p->counts_valid = 0;
    // enqueue helper: false (and rc set, once) when the runtime refused
    auto ok = [&](cudaError_t e, const char* what) {
        if (e == cudaSuccess) return true;
        if (rc == TK_SUCCESS) {
            tk_error_set_detail("process_host: %s -> %s", what, cudaGetErrorString(e));
            rc = TK_ERROR_GPU_MEMORY_COPY_FAILED;
        }
        cudaGetLastError();
        return false;
    };

    // Whole-batch post-process when the logits are device-addressable where they lie.  The caller's d_logits (the
    // _frames entry point) must be complete and visible to the device before the call: they are read on the
    // pipeline's own streams, which nothing orders behind the stream that produced them.
    const float* logits_in_place = d_logits_all;
    bool pack_objectness = false;
    if (h_logits && p->cfg.stage_host_logits != 1) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, h_logits) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
            logits_in_place = static_cast<const float*>(at.devicePointer);
            // a frame or two: one read request per row over PCIe (~42 us per 25200 rows) beats a single-threaded gather
            pack_objectness = p->cfg.stage_host_logits == 0 && (size_t)n * p->cfg.num_proposals >= 65536;
        } else {
            cudaGetLastError();
        }
    }
    if (pack_objectness && (rc = ensure_pack_resources(p)) != TK_SUCCESS) return rc;
    auto post_in_place = [&](const float* d_obj) {
        const tk_error_code_t r = run_kernels(p, nullptr, logits_in_place, n, nullptr, p->d_cands, p->d_counts, p->d_dets, p->d_ndets,
                                              p->s_post, nullptr, d_obj);
        if (r != TK_SUCCESS) { if (rc == TK_SUCCESS) rc = r; return; }
        if (!ok(cudaMemcpyAsync(h_dets, p->d_dets, (size_t)n * det_stride * sizeof(tk_b200_candidate_t), cudaMemcpyDeviceToHost, p->s_post), "detections D2H")) return;
        if (!ok(cudaMemcpyAsync(h_ndets, p->d_ndets, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, p->s_post), "counts D2H")) return;
        if (ok(cudaMemcpyAsync(p->h_cand_counts, p->d_counts, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, p->s_post), "candidate counts D2H"))
            p->counts_valid = n;
    };
    if (logits_in_place && !pack_objectness) post_in_place(nullptr);
    // Frames: when the geometry only ever stages two source rows per output row (4K -> 640: 59 % of the luma rows) and
    // the frames are page-locked, the pre-process kernel's row ring pulls exactly those rows over PCIe in place —
    // fewer bytes cross the link than a copy of the frames would move (tools/exp_pre_zerocopy.py: 32 x 4K I420
    // 8.0 -> 6.1 ms, 4K RGB8 14.5 -> 9.2 ms).  Dense geometries (1080p -> 640 reads every row) keep the chunked copy.
    const void* frames_in_place = nullptr;
    if (h_frames && rc == TK_SUCCESS) {
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, h_frames) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
            const tk_b200_preprocess_batch_params_t b = pre_params(p, at.devicePointer, n, p->d_tensor);
            if (preprocess_reads_rows_sparsely(&b)) frames_in_place = at.devicePointer;
        } else {
            cudaGetLastError();
        }
    }
    if (frames_in_place && rc == TK_SUCCESS) {
        const tk_error_code_t r = run_kernels(p, frames_in_place, nullptr, n, p->d_tensor, nullptr, nullptr, nullptr, nullptr, p->s_compute);
        if (r != TK_SUCCESS) rc = r;
        if (rc == TK_SUCCESS && h_tensor) {
            if (ok(cudaEventRecord(p->ev_computed[0], p->s_compute), "event record") &&
                ok(cudaStreamWaitEvent(p->s_down, p->ev_computed[0], 0), "stream wait"))
                ok(cudaMemcpyAsync(h_tensor, p->d_tensor, (size_t)n * p->tensor_bytes, cudaMemcpyDeviceToHost, p->s_down), "tensor D2H");
        }
        h_frames = nullptr;                       // done: the chunk loop below only has pageable logits left to stage
    }
    const bool chunk_logits = h_logits && !logits_in_place;
    const uint32_t chunks = (h_frames || chunk_logits) ? (n + p->chunk - 1) / p->chunk : 0;
    for (uint32_t k = 0; k < chunks && rc == TK_SUCCESS; ++k) {
        const int s = (int)(k % kSlots);
        const uint32_t f0 = k * p->chunk;
        const uint32_t m = (n - f0) < p->chunk ? (n - f0) : p->chunk;
        // ---- upload
        if (k >= (uint32_t)kSlots && !ok(cudaStreamWaitEvent(p->s_up, p->ev_computed[s], 0), "stream wait")) break;
        if (h_frames && !ok(cudaMemcpyAsync(p->d_frames[s], static_cast<const uint8_t*>(h_frames) + (size_t)f0 * p->frame_bytes,
                                            (size_t)m * p->frame_bytes, cudaMemcpyHostToDevice, p->s_up), "frames H2D")) break;
        if (chunk_logits && !ok(cudaMemcpyAsync(p->d_logits[s], reinterpret_cast<const uint8_t*>(h_logits) + (size_t)f0 * p->logit_bytes,
                                                (size_t)m * p->logit_bytes, cudaMemcpyHostToDevice, p->s_up), "logits H2D")) break;
        if (!ok(cudaEventRecord(p->ev_uploaded[s], p->s_up), "event record")) break;
        // ---- compute
        if (!ok(cudaStreamWaitEvent(p->s_compute, p->ev_uploaded[s], 0), "stream wait")) break;
        const tk_error_code_t r = run_kernels(p, h_frames ? p->d_frames[s] : nullptr, chunk_logits ? p->d_logits[s] : nullptr, m,
                                              p->d_tensor + (size_t)f0 * p->tensor_bytes,
                                              p->d_cands + (size_t)f0 * p->cfg.max_candidates, p->d_counts + f0,
                                              p->d_dets + (size_t)f0 * det_stride, p->d_ndets + f0, p->s_compute);
        if (r != TK_SUCCESS) { rc = r; break; }
        if (!ok(cudaEventRecord(p->ev_computed[s], p->s_compute), "event record")) break;
        // ---- download
        if ((chunk_logits || (h_tensor && h_frames)) && !ok(cudaStreamWaitEvent(p->s_down, p->ev_computed[s], 0), "stream wait")) break;
        if (chunk_logits) {
            if (!ok(cudaMemcpyAsync(h_dets + (size_t)f0 * det_stride, p->d_dets + (size_t)f0 * det_stride,
                                    (size_t)m * det_stride * sizeof(tk_b200_candidate_t), cudaMemcpyDeviceToHost, p->s_down), "detections D2H")) break;
            if (!ok(cudaMemcpyAsync(h_ndets + f0, p->d_ndets + f0, (size_t)m * sizeof(uint32_t), cudaMemcpyDeviceToHost, p->s_down), "counts D2H")) break;
            if (!ok(cudaMemcpyAsync(p->h_cand_counts + f0, p->d_counts + f0, (size_t)m * sizeof(uint32_t), cudaMemcpyDeviceToHost, p->s_down), "candidate counts D2H")) break;
            p->counts_valid = f0 + m;
        }
        if (h_tensor && h_frames &&
            !ok(cudaMemcpyAsync(static_cast<uint8_t*>(h_tensor) + (size_t)f0 * p->tensor_bytes,
                                p->d_tensor + (size_t)f0 * p->tensor_bytes, (size_t)m * p->tensor_bytes,
                                cudaMemcpyDeviceToHost, p->s_down), "tensor D2H")) break;
    }
    if (pack_objectness && rc == TK_SUCCESS) {
        // The frame copies are queued; while the copy engine works through them, the handle's host threads pack the
        // objectness word of every row (4 bytes out of 340) into the page-locked column, which then crosses PCIe as
        // one dense copy instead of one read request per row (profiles/r1_pcie_inplace.md).  Pure data movement: the
        // threshold, like everything else, is applied on the device.
        const size_t P = p->cfg.num_proposals, L = 5 + (size_t)p->cfg.num_classes;
        const size_t rows = (size_t)n * P;
        float* dst = p->h_obj;
        p->pool->parallel_for([=](unsigned part, unsigned parts) {
            const size_t a = rows * part / parts, b = rows * (part + 1) / parts;
            const float* sp = h_logits + a * L + 4;
            for (size_t r = a; r < b; ++r, sp += L) dst[r] = *sp;
        });
        if (ok(cudaMemcpyAsync(p->d_obj, p->h_obj, rows * sizeof(float), cudaMemcpyHostToDevice, p->s_post), "objectness H2D"))
            post_in_place(p->d_obj);
    }
    // Drain, also after an error: no copy stays in flight on the caller's buffers.
    cudaError_t es[4] = {cudaStreamSynchronize(p->s_up), cudaStreamSynchronize(p->s_compute),
                         cudaStreamSynchronize(p->s_down), cudaStreamSynchronize(p->s_post)};
    for (int i = 0; i < 4; ++i) {
        if (es[i] != cudaSuccess && rc == TK_SUCCESS) {
            tk_error_set_detail("pipeline drain: %s", cudaGetErrorString(es[i]));
            rc = TK_ERROR_CUDA_KERNEL_FAILURE;
        }
    }
    cudaGetLastError();
    if (rc == TK_SUCCESS) {
        // rows over the confidence threshold per frame vs the cap the kernels hold (tk_b200_pipeline_candidate_overflow)
        for (uint32_t i = 0; i < p->counts_valid; ++i) {
            const uint32_t c = p->h_cand_counts[i];
            if (c > p->cfg.max_candidates) ++p->over_frames;
            if (c > p->over_max) p->over_max = c;
        }
        if (p->over_frames)
            tk_error_set_detail("process_host: %u frame(s) had more rows over the confidence threshold (up to %u) than "
                                "max_candidates %u; their lists were cut to the first %u rows before NMS",
                                p->over_frames, p->over_max, p->cfg.max_candidates, p->cfg.max_candidates);
    }
    return rc;
}

extern "C" tk_error_code_t tk_b200_pipeline_process_host(tk_b200_pipeline_t* p, const void* h_frames, const float* h_logits,
                                                         uint32_t n, void* h_tensor, tk_b200_candidate_t* h_dets, uint32_t* h_ndets) {
    return process_host_impl(p, h_frames, h_logits, nullptr, n, h_tensor, h_dets, h_ndets);
}

extern "C" tk_error_code_t tk_b200_pipeline_process_host_frames(tk_b200_pipeline_t* p, const void* h_frames, const float* d_logits,
                                                                uint32_t n, void* h_tensor, tk_b200_candidate_t* h_dets, uint32_t* h_ndets) {
    if (!h_frames) return TK_ERROR_INVALID_ARGUMENT;
    return process_host_impl(p, h_frames, nullptr, d_logits, n, h_tensor, h_dets, h_ndets);
}
