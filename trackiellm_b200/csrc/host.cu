// host.cu — entry points that take HOST buffers (the reference's CPU-facing API),
// and tk_cuda_tensor_nms.  Each uploads, runs the same kernels as the device-side
// launchers, and downloads; nothing here computes on the CPU.
//
//   tk_preprocessor_resize_and_normalize_to_chw  src/vision/tk_image_preprocessor.h:49-56
//   tk_classify_dominant_color / _door_state     src/vision/tk_attribute_classifier.h
//   tk_object_detector_free_results              src/vision/tk_object_detector.h:157
//   tk_cuda_tensor_nms                           src/gpu/extensions/cuda/tk_cuda_tensor_ops.h:513-522
//   tk_b200_postprocess_detections (additive)    completes src/vision/tk_object_detector.c:303-368
#include "tk_common.cuh"

#include <stdlib.h>
#include <string.h>

struct tk_gpu_buffer_s { void* d_ptr; size_t size_bytes; int device_id; };   // same layout as dispatch.cu

namespace tkb {

// Per-thread scratch: one stream and grow-only device buffers on the thread's
// current device.  The reference's CPU functions are re-entrant with no global
// state; thread-local scratch keeps that property without a lock.
struct HostScratch {
    int device = -1;
    cudaStream_t stream = nullptr;
    void* buf[4] = {nullptr, nullptr, nullptr, nullptr};
    size_t cap[4] = {0, 0, 0, 0};
    // Frees the thread's stream and device buffers (thread exit, or tk_b200_thread_release): a short-lived thread that
    // calls a host entry point must not leave device memory behind.  At process teardown the runtime may already be
    // gone; every call here tolerates that.
    void release() {
        if (device < 0) return;
        int cur = -1;
        const bool have_cur = cudaGetDevice(&cur) == cudaSuccess;
        if (cudaSetDevice(device) == cudaSuccess) {
            if (stream) cudaStreamSynchronize(stream);
            for (int i = 0; i < 4; ++i) { if (buf[i]) cudaFree(buf[i]); buf[i] = nullptr; cap[i] = 0; }
            if (stream) cudaStreamDestroy(stream);
            if (have_cur && cur >= 0) cudaSetDevice(cur);
        }
        cudaGetLastError();
        stream = nullptr; device = -1;
    }
    ~HostScratch() { release(); }
};
static thread_local HostScratch g_hs;

static tk_error_code_t scratch_prepare(HostScratch** out) {
    int dev = -1;
    if (!device_info(&dev)) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    HostScratch& s = g_hs;
    if (s.device != dev) {
        // device changed under this thread: drop the old resources (best effort)
        if (s.device >= 0) {
            int cur = dev; cudaSetDevice(s.device);
            for (int i = 0; i < 4; ++i) { if (s.buf[i]) cudaFree(s.buf[i]); s.buf[i] = nullptr; s.cap[i] = 0; }
            if (s.stream) cudaStreamDestroy(s.stream);
            cudaSetDevice(cur); cudaGetLastError();
        }
        s.stream = nullptr; s.device = dev;
        TK_CUDA_TRY(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking), TK_ERROR_GPU_STREAM_CREATION_FAILED);
    }
    *out = &s;
    return TK_SUCCESS;
}

static tk_error_code_t scratch_reserve(HostScratch* s, int slot, size_t bytes) {
    if (s->cap[slot] >= bytes) return TK_SUCCESS;
    if (s->buf[slot]) { cudaStreamSynchronize(s->stream); cudaFree(s->buf[slot]); s->buf[slot] = nullptr; s->cap[slot] = 0; }
    size_t want = bytes + bytes / 4 + 256;
    if (cudaMalloc(&s->buf[slot], want) != cudaSuccess) {
        tk_error_set_detail("cudaMalloc(%zu) failed: %s", want, cudaGetErrorString(cudaGetLastError()));
        return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
    }
    s->cap[slot] = want;
    return TK_SUCCESS;
}

// Candidate list from (boxes f32 [K][4] xywh, scores f32 [K]): score >= thr
// (predicate of tk_cuda_tensor_ops.cu:476), box truncated like
// tk_object_detector.c:347-350, ascending box index, one CTA.
__global__ void __launch_bounds__(1024) boxes_to_candidates_kernel(const float* __restrict__ boxes, const float* __restrict__ scores,
                                                                   uint32_t K, float thr, tk_b200_candidate_t* __restrict__ out,
                                                                   uint32_t cap, uint32_t* __restrict__ count) {
    __shared__ uint32_t wtot[32];
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    uint32_t run = 0;
    for (uint32_t base = 0; base < K; base += 1024) {
        const uint32_t i = base + tid;
        const float sc = i < K ? scores[i] : 0.0f;
        const bool pass = i < K && sc >= thr;
        const uint32_t bal = __ballot_sync(0xffffffffu, pass);
        if (lane == 0) wtot[warp] = __popc(bal);
        __syncthreads();
        uint32_t before = 0, tot = 0;
        for (uint32_t w = 0; w < 32; ++w) { uint32_t t = wtot[w]; before += w < warp ? t : 0u; tot += t; }
        const uint32_t o = run + before + __popc(bal & ((1u << lane) - 1u));
        if (pass && o < cap) {
            const float4 b = *reinterpret_cast<const float4*>(boxes + 4 * (size_t)i);
            tk_b200_candidate_t c;
            c.index = (int32_t)i;
            c.bbox.x = __float2int_rz(b.x); c.bbox.y = __float2int_rz(b.y);
            c.bbox.w = __float2int_rz(b.z); c.bbox.h = __float2int_rz(b.w);
            c.confidence = sc; c.class_id = 0;
            out[o] = c;
        }
        run += tot;
        __syncthreads();
    }
    if (tid == 0) *count = run;
}

__global__ void candidates_to_indices_kernel(const tk_b200_candidate_t* __restrict__ dets, const uint32_t* __restrict__ n,
                                             int32_t* __restrict__ indices, uint32_t md, uint32_t max_det) {
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < max_det;
         i += (unsigned long long)gridDim.x * blockDim.x)
        indices[i] = (i < md && i < *n) ? dets[i].index : -1;
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_cuda_tensor_nms(tk_cuda_dispatcher_t* dispatcher, const tk_tensor_descriptor_t* boxes,
                                              const tk_tensor_descriptor_t* scores, tk_gpu_buffer_t indices,
                                              uint32_t* num_detections, uint32_t max_detections, float iou_threshold,
                                              float score_threshold) {
    if (!dispatcher || !boxes || !scores || !indices || !num_detections) return TK_ERROR_INVALID_ARGUMENT;
    if (!boxes->buffer || !scores->buffer || !boxes->buffer->d_ptr || !scores->buffer->d_ptr || !indices->d_ptr)
        return TK_ERROR_GPU_INVALID_BUFFER;
    if (boxes->type != TK_TENSOR_TYPE_F32 || scores->type != TK_TENSOR_TYPE_F32) return TK_ERROR_INVALID_INPUT_TENSOR;
    if (boxes->dimensions < 2) return TK_ERROR_INVALID_DIMENSIONS;
    const uint32_t K = boxes->shape[1];                                  // tk_cuda_tensor_ops.cu:883
    *num_detections = 0;
    if (K == 0 || max_detections == 0) return TK_SUCCESS;
    if ((size_t)K * 16 > boxes->buffer->size_bytes || (size_t)K * 4 > scores->buffer->size_bytes ||
        (size_t)max_detections * 4 > indices->size_bytes)
        return TK_ERROR_GPU_BUFFER_SIZE_MISMATCH;
    const int dev = tk_b200_dispatch_device(dispatcher);
    int prev = -1; cudaGetDevice(&prev);
    if (cudaSetDevice(dev) != cudaSuccess) { cudaGetLastError(); return TK_ERROR_GPU_DEVICE_SET_FAILED; }
    cudaStream_t st = tk_b200_dispatch_stream(dispatcher);
    const uint32_t cap = K;                                    // every box is a candidate (above 4096: the global NMS path)
    const uint32_t md = max_detections < cap ? max_detections : cap;
    tk_error_code_t rc = TK_SUCCESS;
    uint8_t* scratch = nullptr;
    const size_t cand_bytes = (size_t)cap * sizeof(tk_b200_candidate_t), det_bytes = (size_t)md * sizeof(tk_b200_candidate_t);
    const size_t off_det = (cand_bytes + 255) & ~(size_t)255, off_cnt = (off_det + det_bytes + 255) & ~(size_t)255;
    if (cudaMallocAsync(reinterpret_cast<void**>(&scratch), off_cnt + 256, st) != cudaSuccess) {
        cudaGetLastError(); rc = TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
    }
    uint32_t host_n = 0;
    if (rc == TK_SUCCESS) {
        tk_b200_candidate_t* d_c = reinterpret_cast<tk_b200_candidate_t*>(scratch);
        tk_b200_candidate_t* d_d = reinterpret_cast<tk_b200_candidate_t*>(scratch + off_det);
        uint32_t* d_cnt = reinterpret_cast<uint32_t*>(scratch + off_cnt);
        boxes_to_candidates_kernel<<<1, 1024, 0, st>>>(static_cast<const float*>(boxes->buffer->d_ptr),
                                                       static_cast<const float*>(scores->buffer->d_ptr), K, score_threshold,
                                                       d_c, cap, d_cnt);
        rc = launch_status("boxes_to_candidates_kernel");
        if (rc == TK_SUCCESS) {
            tk_b200_nms_params_t np{};
            np.d_candidates = d_c; np.d_counts = d_cnt; np.num_frames = 1; np.max_candidates = cap;
            np.iou_threshold = iou_threshold; np.max_detections = md; np.class_agnostic = 1;
            np.d_detections = d_d; np.d_num_detections = d_cnt + 1;
            rc = tk_kernels_nms_classwise(&np, st);
        }
        if (rc == TK_SUCCESS) {
            // indices[0, n) = kept boxes, [n, max_detections) = -1 (the whole buffer the caller sized is defined)
            uint32_t blocks = max_detections / 256 + (max_detections % 256 ? 1u : 0u);
            if (blocks > 1024u) blocks = 1024u;                 // grid-stride loop inside
            candidates_to_indices_kernel<<<blocks, 256, 0, st>>>(d_d, d_cnt + 1, static_cast<int32_t*>(indices->d_ptr), md, max_detections);
            rc = launch_status("candidates_to_indices_kernel");
        }
        if (rc == TK_SUCCESS && cudaMemcpyAsync(&host_n, d_cnt + 1, 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) rc = TK_ERROR_GPU_MEMORY_COPY_FAILED;
        cudaFreeAsync(scratch, st);
        if (cudaStreamSynchronize(st) != cudaSuccess) { if (rc == TK_SUCCESS) rc = TK_ERROR_GPU_SYNCHRONIZATION_FAILED; }
        cudaGetLastError();
    }
    if (rc == TK_SUCCESS) *num_detections = host_n;
    if (prev >= 0) cudaSetDevice(prev);
    return rc;
}

// ---------------------------------------------------------------------------
// src/vision/tk_image_preprocessor.h:49-56 — HOST frame in, HOST fp32 CHW out.
// Argument checks follow src/vision/tk_image_preprocessor.c:29-31.
// ---------------------------------------------------------------------------
extern "C" tk_error_code_t tk_preprocessor_resize_and_normalize_to_chw(const tk_video_frame_t* frame, float* out_tensor,
                                                                       uint32_t target_width, uint32_t target_height,
                                                                       const float mean[3], const float std_dev[3]) {
    if (!frame || !frame->data || !out_tensor || !mean || !std_dev || target_width == 0 || target_height == 0)
        return TK_ERROR_INVALID_ARGUMENT;
    if (frame->width < 2 || frame->height < 2) return TK_ERROR_INVALID_ARGUMENT;
    uint32_t bpp; size_t in_bytes;
    switch (frame->format) {
        case TK_PIXEL_FORMAT_RGB8: case TK_PIXEL_FORMAT_BGR8: bpp = 3; break;
        case TK_PIXEL_FORMAT_RGBA8: bpp = 4; break;
        case TK_PIXEL_FORMAT_I420: bpp = 1; break;
        default: return TK_ERROR_INVALID_ARGUMENT;
    }
    const uint32_t stride = frame->stride ? frame->stride : frame->width * bpp;
    if (stride < frame->width * bpp) return TK_ERROR_INVALID_ARGUMENT;
    in_bytes = (size_t)stride * frame->height;
    if (frame->format == TK_PIXEL_FORMAT_I420) in_bytes += 2 * (size_t)(stride / 2) * (frame->height / 2);
    const size_t out_bytes = (size_t)target_width * target_height * 3 * sizeof(float);
    HostScratch* s;
    tk_error_code_t rc = scratch_prepare(&s);
    if (rc != TK_SUCCESS) return rc;
    if ((rc = scratch_reserve(s, 0, in_bytes + 64)) != TK_SUCCESS) return rc;      // +64: the row ring's 16-byte copies may overrun the last row
    if ((rc = scratch_reserve(s, 1, out_bytes)) != TK_SUCCESS) return rc;
    TK_CUDA_TRY(cudaMemcpyAsync(s->buf[0], frame->data, in_bytes, cudaMemcpyHostToDevice, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    tk_b200_preprocess_batch_params_t b{};
    b.d_input = s->buf[0]; b.num_frames = 1;
    b.input_width = frame->width; b.input_height = frame->height; b.input_stride_bytes = stride;
    b.input_format = frame->format;
    b.d_output = s->buf[1]; b.output_width = target_width; b.output_height = target_height;
    b.output_type = TK_B200_OUT_F32; b.resize_mode = TK_B200_RESIZE_STRETCH; b.arith = TK_B200_ARITH_EXACT;
    b.mean = TkFloat3{mean[0], mean[1], mean[2]};
    b.std_dev = TkFloat3{std_dev[0], std_dev[1], std_dev[2]};
    if ((rc = tk_b200_preprocess_batch(&b, s->stream)) != TK_SUCCESS) return rc;
    TK_CUDA_TRY(cudaMemcpyAsync(out_tensor, s->buf[1], out_bytes, cudaMemcpyDeviceToHost, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaStreamSynchronize(s->stream), TK_ERROR_CUDA_KERNEL_FAILURE);
    return TK_SUCCESS;
}

// Uploads only the frame rows the box can touch.
static tk_error_code_t classify_host(const tk_video_frame_t* frame, const tk_rect_t* bbox, bool door, uint32_t* result) {
    if (frame->format != TK_PIXEL_FORMAT_RGB8) return TK_ERROR_INVALID_ARGUMENT;
    if (frame->width == 0 || frame->height == 0 || !frame->data) return TK_ERROR_INVALID_ARGUMENT;
    const uint32_t stride = frame->stride ? frame->stride : frame->width * 3;
    long long y0 = bbox->y, y1 = (long long)bbox->y + bbox->h;
    if (y0 < 0) y0 = 0;
    if (y1 > (long long)frame->height) y1 = frame->height;
    HostScratch* s;
    tk_error_code_t rc = scratch_prepare(&s);
    if (rc != TK_SUCCESS) return rc;
    const size_t rows = y1 > y0 ? (size_t)(y1 - y0) : 0;
    const size_t bytes = rows * stride;
    if ((rc = scratch_reserve(s, 0, bytes + 64)) != TK_SUCCESS) return rc;
    if ((rc = scratch_reserve(s, 2, 256)) != TK_SUCCESS) return rc;
    if (bytes) TK_CUDA_TRY(cudaMemcpyAsync(s->buf[0], frame->data + (size_t)y0 * stride, bytes, cudaMemcpyHostToDevice, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    tk_rect_t local = *bbox;
    local.y = (int)((long long)bbox->y - y0);          // box in the coordinates of the uploaded row window
    uint8_t* sc = static_cast<uint8_t*>(s->buf[2]);
    TK_CUDA_TRY(cudaMemcpyAsync(sc, &local, sizeof(local), cudaMemcpyHostToDevice, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    const uint32_t win_h = rows ? (uint32_t)rows : 1u;
    if (door) {
        tk_b200_door_params_t p{};
        p.d_frame = static_cast<const uint8_t*>(s->buf[0]); p.width = frame->width; p.height = win_h; p.stride_bytes = stride;
        p.d_boxes = reinterpret_cast<const tk_rect_t*>(sc); p.num_boxes = 1;
        p.d_states = reinterpret_cast<uint32_t*>(sc + 64); p.d_edge_pixels = nullptr;
        rc = tk_kernels_door_state(&p, s->stream);
    } else {
        tk_b200_color_params_t p{};
        p.d_frame = static_cast<const uint8_t*>(s->buf[0]); p.width = frame->width; p.height = win_h; p.stride_bytes = stride;
        p.d_boxes = reinterpret_cast<const tk_rect_t*>(sc); p.num_boxes = 1;
        p.d_color_ids = reinterpret_cast<uint32_t*>(sc + 64); p.d_bins = nullptr;
        rc = tk_kernels_dominant_color(&p, s->stream);
    }
    if (rc != TK_SUCCESS) return rc;
    TK_CUDA_TRY(cudaMemcpyAsync(result, sc + 64, 4, cudaMemcpyDeviceToHost, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaStreamSynchronize(s->stream), TK_ERROR_CUDA_KERNEL_FAILURE);
    return TK_SUCCESS;
}

extern "C" tk_error_code_t tk_classify_dominant_color(const tk_video_frame_t* frame, const tk_rect_t* bbox, char** out_color_name) {
    if (!frame || !bbox || !out_color_name) return TK_ERROR_INVALID_ARGUMENT;    // tk_attribute_classifier.c:45-47
    uint32_t id = 0;
    tk_error_code_t rc = classify_host(frame, bbox, false, &id);
    if (rc != TK_SUCCESS) return rc;
    *out_color_name = strdup(tk_b200_color_name(id));
    return *out_color_name ? TK_SUCCESS : TK_ERROR_OUT_OF_MEMORY;
}

extern "C" tk_error_code_t tk_classify_door_state(const tk_video_frame_t* frame, const tk_rect_t* bbox, char** out_state_name) {
    if (!frame || !bbox || !out_state_name) return TK_ERROR_INVALID_ARGUMENT;    // tk_attribute_classifier.c:105-107
    uint32_t st = 0;
    tk_error_code_t rc = classify_host(frame, bbox, true, &st);
    if (rc != TK_SUCCESS) return rc;
    *out_state_name = strdup(st ? "closed" : "open");
    return *out_state_name ? TK_SUCCESS : TK_ERROR_OUT_OF_MEMORY;
}

// ---------------------------------------------------------------------------
// postprocess_detections completed (src/vision/tk_object_detector.c:303-368):
// HOST logits [P][5+C] -> malloc'd tk_detection_result_t[] (:359-361 is the copy
// the reference elides), boxes restored to source-frame px (:365).  Decode and
// NMS run on the GPU; the host only marshals the kept records into the
// reference's result struct.
// ---------------------------------------------------------------------------
extern "C" tk_error_code_t tk_b200_postprocess_detections(
    const float* h_logits, uint32_t P, uint32_t C, tk_b200_activation_e act, float conf_thr, float iou_thr,
    uint32_t max_det, const char* const* class_labels, size_t class_count, uint32_t frame_w, uint32_t frame_h,
    uint32_t in_w, uint32_t in_h, tk_detection_result_t** out_results, size_t* out_count) {
    if (!h_logits || !out_results || !out_count) return TK_ERROR_INVALID_ARGUMENT;
    *out_results = nullptr; *out_count = 0;
    if (P == 0) return TK_ERROR_INVALID_DIMENSIONS;
    if (max_det == 0) max_det = 500;                       // TK_OBJECT_DETECTOR_MAX_DETECTIONS, :39
    const uint32_t cap = 4096;
    const size_t logit_bytes = (size_t)P * (5 + (size_t)C) * sizeof(float);
    const size_t cand_bytes = (size_t)cap * sizeof(tk_b200_candidate_t);
    const size_t det_bytes = (size_t)max_det * sizeof(tk_b200_candidate_t);
    HostScratch* s;
    tk_error_code_t rc = scratch_prepare(&s);
    if (rc != TK_SUCCESS) return rc;
    if ((rc = scratch_reserve(s, 0, logit_bytes)) != TK_SUCCESS) return rc;
    if ((rc = scratch_reserve(s, 1, cand_bytes + det_bytes + 512)) != TK_SUCCESS) return rc;
    uint8_t* base = static_cast<uint8_t*>(s->buf[1]);
    tk_b200_candidate_t* d_c = reinterpret_cast<tk_b200_candidate_t*>(base);
    const size_t off_d = (cand_bytes + 255) & ~(size_t)255;
    tk_b200_candidate_t* d_d = reinterpret_cast<tk_b200_candidate_t*>(base + off_d);
    uint32_t* d_n = reinterpret_cast<uint32_t*>(base + ((off_d + det_bytes + 255) & ~(size_t)255));
    TK_CUDA_TRY(cudaMemcpyAsync(s->buf[0], h_logits, logit_bytes, cudaMemcpyHostToDevice, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    tk_b200_decode_params_t d{};
    d.d_logits = static_cast<const float*>(s->buf[0]); d.num_frames = 1; d.num_proposals = P; d.num_classes = C;
    d.confidence_threshold = conf_thr; d.activation = act; d.d_candidates = d_c; d.d_counts = d_n; d.max_candidates = cap;
    if ((rc = tk_kernels_decode_detections(&d, s->stream)) != TK_SUCCESS) return rc;
    tk_b200_nms_params_t m{};
    m.d_candidates = d_c; m.d_counts = d_n; m.num_frames = 1; m.max_candidates = cap; m.iou_threshold = iou_thr;
    m.max_detections = max_det; m.class_agnostic = 0; m.d_detections = d_d; m.d_num_detections = d_n + 1;
    if (frame_w && frame_h && in_w && in_h) {
        m.restore_src_width = frame_w; m.restore_src_height = frame_h; m.restore_in_width = in_w; m.restore_in_height = in_h;
    }
    if ((rc = tk_kernels_nms_classwise(&m, s->stream)) != TK_SUCCESS) return rc;
    uint32_t cnt[2] = {0, 0};                                  // rows over the threshold, kept detections
    TK_CUDA_TRY(cudaMemcpyAsync(cnt, d_n, 8, cudaMemcpyDeviceToHost, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaStreamSynchronize(s->stream), TK_ERROR_CUDA_KERNEL_FAILURE);
    if (cnt[0] > cap) {
        // more candidates than the capped kernels hold: the reference keeps them all (:339-355) — again, uncapped
        const uint32_t restore[4] = {frame_w, frame_h, in_w, in_h};
        const bool any = frame_w && frame_h && in_w && in_h;
        if ((rc = postprocess_frame_uncapped(static_cast<const float*>(s->buf[0]), P, C, act, conf_thr, iou_thr, max_det, 0,
                                             any ? restore : nullptr, d_d, d_n + 1, s->stream)) != TK_SUCCESS) return rc;
        TK_CUDA_TRY(cudaMemcpyAsync(cnt, d_n, 8, cudaMemcpyDeviceToHost, s->stream), TK_ERROR_GPU_MEMORY_COPY_FAILED);
        TK_CUDA_TRY(cudaStreamSynchronize(s->stream), TK_ERROR_CUDA_KERNEL_FAILURE);
    }
    const uint32_t n = cnt[1];
    if (n == 0) return TK_SUCCESS;
    tk_b200_candidate_t* tmp = static_cast<tk_b200_candidate_t*>(malloc((size_t)n * sizeof(tk_b200_candidate_t)));
    tk_detection_result_t* res = static_cast<tk_detection_result_t*>(calloc(n, sizeof(tk_detection_result_t)));
    if (!tmp || !res) { free(tmp); free(res); return TK_ERROR_OUT_OF_MEMORY; }
    if (cudaMemcpy(tmp, d_d, (size_t)n * sizeof(tk_b200_candidate_t), cudaMemcpyDeviceToHost) != cudaSuccess) {
        cudaGetLastError(); free(tmp); free(res); return TK_ERROR_GPU_MEMORY_COPY_FAILED;
    }
    for (uint32_t i = 0; i < n; ++i) {
        res[i].class_id = tmp[i].class_id;
        res[i].label = (class_labels && tmp[i].class_id < class_count) ? class_labels[tmp[i].class_id] : nullptr;
        res[i].confidence = tmp[i].confidence;
        res[i].bbox = tmp[i].bbox;
    }
    free(tmp);
    *out_results = res; *out_count = n;
    return TK_SUCCESS;
}

// src/vision/tk_object_detector.c:205-210
extern "C" void tk_object_detector_free_results(tk_detection_result_t** results) {
    if (results && *results) { free(*results); *results = nullptr; }
}

// Drops the calling thread's scratch (stream + device buffers of the host-pointer entry points) now instead of at
// thread exit.
extern "C" void tk_b200_thread_release(void) { tkb::g_hs.release(); }

// Pinned host memory so the batched host entry points copy by DMA.
extern "C" tk_error_code_t tk_b200_host_alloc(void** out, size_t bytes) {
    if (!out || bytes == 0) return TK_ERROR_INVALID_ARGUMENT;
    *out = nullptr;
    if (!device_info()) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    if (cudaHostAlloc(out, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return TK_ERROR_OUT_OF_MEMORY; }
    return TK_SUCCESS;
}
extern "C" void tk_b200_host_free(void* p) { if (p) { cudaFreeHost(p); cudaGetLastError(); } }
