// dispatch.cu — GPU dispatch layer: device/stream/buffer handles, async copies,
// workflow wrappers, the depth kernels and tk_cuda_tensor_nms.
//
// Replaces  tk_cuda_dispatch_* (src/gpu/cuda/tk_cuda_dispatch.h:72-171, impl
//             src/gpu/cuda/tk_cuda_dispatch.cu:163-585),
//           tk_kernels_postprocess_depth_map / tk_kernels_depth_to_point_cloud
//             (src/gpu/cuda/tk_cuda_kernels.cu:172-249, :396-500),
//           tk_cuda_tensor_nms (src/gpu/extensions/cuda/tk_cuda_tensor_ops.cu:858-889).
//
// Differences from the reference, all deliberate:
//   * upload -> compute -> download are ordered by events (the reference creates
//     three non-blocking streams and never links them, tk_cuda_dispatch.cu:107-137,
//     so its documented flow races);
//   * destroy does not cudaDeviceReset() (tk_cuda_dispatch.cu:245-247 would kill
//     sibling dispatchers in a multi-GPU process);
//   * every entry point selects the dispatcher's device for the calling thread
//     and restores the previous one.
#include "tk_common.cuh"

#include <stdlib.h>
#include <string.h>
#include <new>

struct tk_gpu_buffer_s {
    void*  d_ptr;
    size_t size_bytes;
    int    device_id;
};

struct tk_cuda_dispatcher_s {
    int device_id;
    cudaStream_t compute, upload, download;
    cudaEvent_t ev_upload, ev_compute;
};

namespace tkb {

struct DeviceScope {
    int prev = -1; bool ok = false;
    explicit DeviceScope(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
        ok = cudaSetDevice(dev) == cudaSuccess;
        if (!ok) cudaGetLastError();
    }
    ~DeviceScope() { if (prev >= 0) cudaSetDevice(prev); }
};

// ---- depth kernels ("next" row of SURVEY §8f; same boundary header) ---------
// metric = raw * scale + shift, the C expression of tk_cuda_kernels.cu:197 with
// one rounding per operation.
__global__ void __launch_bounds__(256) depth_scale_kernel(const float* __restrict__ in, float* __restrict__ out, size_t n,
                                                          float scale, float shift, int vec) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t step = (size_t)gridDim.x * blockDim.x;
    if (vec) {
        const size_t n4 = n >> 2;
        auto conv = [&](float4 v) {
            v.x = __fadd_rn(__fmul_rn(v.x, scale), shift); v.y = __fadd_rn(__fmul_rn(v.y, scale), shift);
            v.z = __fadd_rn(__fmul_rn(v.z, scale), shift); v.w = __fadd_rn(__fmul_rn(v.w, scale), shift);
            return v;
        };
        // four 128-bit loads in flight per thread before the first store: enough bytes outstanding per SM for HBM3e
        for (; i + 3 * step < n4; i += 4 * step) {
            const float4 a = __ldcs(reinterpret_cast<const float4*>(in) + i);
            const float4 b = __ldcs(reinterpret_cast<const float4*>(in) + i + step);
            const float4 c = __ldcs(reinterpret_cast<const float4*>(in) + i + 2 * step);
            const float4 d = __ldcs(reinterpret_cast<const float4*>(in) + i + 3 * step);
            __stcs(reinterpret_cast<float4*>(out) + i, conv(a));
            __stcs(reinterpret_cast<float4*>(out) + i + step, conv(b));
            __stcs(reinterpret_cast<float4*>(out) + i + 2 * step, conv(c));
            __stcs(reinterpret_cast<float4*>(out) + i + 3 * step, conv(d));
        }
        for (; i < n4; i += step) __stcs(reinterpret_cast<float4*>(out) + i, conv(__ldcs(reinterpret_cast<const float4*>(in) + i)));
        i = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    }
    for (; i < n; i += step) out[i] = __fadd_rn(__fmul_rn(in[i], scale), shift);
}

// tk_cuda_kernels.cu:201-249: x = (u - cx) * d / fx, y = (v - cy) * d / fy, z = d;
// zero when d <= 0 or d > 1000.  Each thread makes 4 points = 3 float4 stores.
__device__ __forceinline__ void unproject(float d, uint32_t u, uint32_t v, float fx, float fy, float cx, float cy, float* o) {
    if (d <= 0.0f || d > 1000.0f) { o[0] = o[1] = o[2] = 0.0f; return; }   // a NaN depth falls through, as in the reference
    o[0] = __fdiv_rn(__fmul_rn(__fsub_rn((float)u, cx), d), fx);
    o[1] = __fdiv_rn(__fmul_rn(__fsub_rn((float)v, cy), d), fy);
    o[2] = d;
}
__global__ void __launch_bounds__(256) depth_points_kernel(const float* __restrict__ depth, float* __restrict__ pts, uint32_t W,
                                                           size_t n, float fx, float fy, float cx, float cy, int vec) {
    const size_t step = (size_t)gridDim.x * blockDim.x;
    size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (vec) {
        const size_t n4 = n >> 2;
        auto emit = [&](size_t qq, const float4 d4) {
            const float ds[4] = {d4.x, d4.y, d4.z, d4.w};
            float o[12];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const size_t idx = qq * 4 + j;
                unproject(ds[j], (uint32_t)(idx % W), (uint32_t)(idx / W), fx, fy, cx, cy, o + 3 * j);
            }
            float4* dst = reinterpret_cast<float4*>(pts) + qq * 3;
            __stcs(dst + 0, make_float4(o[0], o[1], o[2], o[3]));
            __stcs(dst + 1, make_float4(o[4], o[5], o[6], o[7]));
            __stcs(dst + 2, make_float4(o[8], o[9], o[10], o[11]));
        };
        for (; q < n4; q += step) emit(q, __ldcs(reinterpret_cast<const float4*>(depth) + q));
        q = (n4 << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    }
    for (; q < n; q += step) unproject(depth[q], (uint32_t)(q % W), (uint32_t)(q / W), fx, fy, cx, cy, pts + 3 * q);
}

static uint32_t stream_grid(const DeviceInfo* dev, size_t work_items, int threads) {
    size_t g = (work_items + threads - 1) / threads;
    const size_t cap = (size_t)dev->sm_count * 8;
    return (uint32_t)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace tkb

using namespace tkb;

extern "C" tk_error_code_t tk_kernels_postprocess_depth_map(const tk_postprocess_depth_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_raw_depth_map || !p->d_metric_depth_map) return TK_ERROR_INVALID_ARGUMENT;
    if (p->width == 0 || p->height == 0) return TK_ERROR_INVALID_DIMENSIONS;
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    const size_t n = (size_t)p->width * p->height;
    const int vec = ((reinterpret_cast<uintptr_t>(p->d_raw_depth_map) | reinterpret_cast<uintptr_t>(p->d_metric_depth_map)) & 15) == 0;
    depth_scale_kernel<<<stream_grid(dev, vec ? n / 4 + 1 : n, 256), 256, 0, stream>>>(
        static_cast<const float*>(p->d_raw_depth_map), static_cast<float*>(p->d_metric_depth_map), n, p->scale, p->shift, vec);
    return launch_status("depth_scale_kernel");
}

extern "C" tk_error_code_t tk_kernels_depth_to_point_cloud(const tk_depth_to_points_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_metric_depth_map || !p->d_point_cloud) return TK_ERROR_INVALID_ARGUMENT;
    if (p->width == 0 || p->height == 0) return TK_ERROR_INVALID_DIMENSIONS;
    if (p->fx <= 0.0f || p->fy <= 0.0f) return TK_ERROR_INVALID_CAMERA_PARAMETERS;      // tk_cuda_kernels.cu:470-472
    const DeviceInfo* dev = device_info();
    if (!dev) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    const size_t n = (size_t)p->width * p->height;
    const int vec = ((reinterpret_cast<uintptr_t>(p->d_metric_depth_map) | reinterpret_cast<uintptr_t>(p->d_point_cloud)) & 15) == 0;
    depth_points_kernel<<<stream_grid(dev, vec ? n / 4 + 1 : n, 256), 256, 0, stream>>>(
        static_cast<const float*>(p->d_metric_depth_map), static_cast<float*>(p->d_point_cloud), p->width, n,
        p->fx, p->fy, p->cx, p->cy, vec);
    return launch_status("depth_points_kernel");
}

// ---------------------------------------------------------------------------
// Dispatcher lifecycle / memory / sync  (src/gpu/cuda/tk_cuda_dispatch.h:72-141)
// ---------------------------------------------------------------------------
extern "C" tk_error_code_t tk_cuda_dispatch_create(tk_cuda_dispatcher_t** out, const tk_cuda_dispatcher_config_t* cfg) {
    if (!out || !cfg) return TK_ERROR_INVALID_ARGUMENT;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return TK_ERROR_GPU_DEVICE_ENUMERATION_FAILED; }
    if (count == 0 || cfg->device_id < 0 || cfg->device_id >= count) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    DeviceScope scope(cfg->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    if (!device_info()) return TK_ERROR_GPU_DEVICE_UNSUPPORTED;       // sm_100a code only
    tk_cuda_dispatcher_t* d = new (std::nothrow) tk_cuda_dispatcher_t();
    if (!d) return TK_ERROR_OUT_OF_MEMORY;
    d->device_id = cfg->device_id;
    bool ok = cudaStreamCreateWithFlags(&d->compute, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&d->upload, cudaStreamNonBlocking) == cudaSuccess &&
              cudaStreamCreateWithFlags(&d->download, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreateWithFlags(&d->ev_upload, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&d->ev_compute, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        cudaGetLastError();
        tk_cuda_dispatch_destroy(&d);
        return TK_ERROR_GPU_STREAM_CREATION_FAILED;
    }
    *out = d;
    return TK_SUCCESS;
}

extern "C" void tk_cuda_dispatch_destroy(tk_cuda_dispatcher_t** pd) {
    if (!pd || !*pd) return;
    tk_cuda_dispatcher_t* d = *pd;
    {
        DeviceScope scope(d->device_id);
        if (d->compute) { cudaStreamSynchronize(d->compute); cudaStreamDestroy(d->compute); }
        if (d->upload) { cudaStreamSynchronize(d->upload); cudaStreamDestroy(d->upload); }
        if (d->download) { cudaStreamSynchronize(d->download); cudaStreamDestroy(d->download); }
        if (d->ev_upload) cudaEventDestroy(d->ev_upload);
        if (d->ev_compute) cudaEventDestroy(d->ev_compute);
        cudaGetLastError();
    }
    delete d;
    *pd = nullptr;
}

extern "C" tk_error_code_t tk_cuda_dispatch_malloc(tk_cuda_dispatcher_t* d, tk_gpu_buffer_t* out, size_t size) {
    if (!d || !out || size == 0) return TK_ERROR_INVALID_ARGUMENT;
    *out = nullptr;
    DeviceScope scope(d->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    tk_gpu_buffer_s* b = static_cast<tk_gpu_buffer_s*>(malloc(sizeof(tk_gpu_buffer_s)));
    if (!b) return TK_ERROR_OUT_OF_MEMORY;
    if (cudaMalloc(&b->d_ptr, size) != cudaSuccess) {
        tk_error_set_detail("cudaMalloc(%zu) failed: %s", size, cudaGetErrorString(cudaGetLastError()));
        free(b);
        return TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED;
    }
    b->size_bytes = size; b->device_id = d->device_id;
    *out = b;
    return TK_SUCCESS;
}

extern "C" void tk_cuda_dispatch_free(tk_cuda_dispatcher_t* d, tk_gpu_buffer_t* buf) {
    if (!d || !buf || !*buf) return;
    DeviceScope scope((*buf)->device_id);
    cudaFree((*buf)->d_ptr);
    cudaGetLastError();
    free(*buf);
    *buf = nullptr;
}

extern "C" tk_error_code_t tk_cuda_dispatch_upload_async(tk_cuda_dispatcher_t* d, tk_gpu_buffer_t dst, const void* src, size_t size) {
    if (!d || !dst || !src) return TK_ERROR_INVALID_ARGUMENT;
    if (!dst->d_ptr) return TK_ERROR_GPU_INVALID_BUFFER;
    if (size > dst->size_bytes) return TK_ERROR_GPU_BUFFER_SIZE_MISMATCH;
    DeviceScope scope(d->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    // The copy may not overtake compute still reading the destination.
    TK_CUDA_TRY(cudaEventRecord(d->ev_compute, d->compute), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaStreamWaitEvent(d->upload, d->ev_compute, 0), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaMemcpyAsync(dst->d_ptr, src, size, cudaMemcpyHostToDevice, d->upload), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    // Anything enqueued on the compute stream from now on sees the uploaded bytes.
    TK_CUDA_TRY(cudaEventRecord(d->ev_upload, d->upload), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaStreamWaitEvent(d->compute, d->ev_upload, 0), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    return TK_SUCCESS;
}

extern "C" tk_error_code_t tk_cuda_dispatch_download_async(tk_cuda_dispatcher_t* d, void* dst, tk_gpu_buffer_t src, size_t size) {
    if (!d || !dst || !src) return TK_ERROR_INVALID_ARGUMENT;
    if (!src->d_ptr) return TK_ERROR_GPU_INVALID_BUFFER;
    if (size > src->size_bytes) return TK_ERROR_GPU_BUFFER_SIZE_MISMATCH;
    DeviceScope scope(d->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    TK_CUDA_TRY(cudaEventRecord(d->ev_compute, d->compute), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaStreamWaitEvent(d->download, d->ev_compute, 0), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    TK_CUDA_TRY(cudaMemcpyAsync(dst, src->d_ptr, size, cudaMemcpyDeviceToHost, d->download), TK_ERROR_GPU_MEMORY_COPY_FAILED);
    return TK_SUCCESS;
}

// The reference waits on the default stream only (tk_cuda_dispatch.cu:475); here
// all three streams are drained so a returned download is complete.
extern "C" tk_error_code_t tk_cuda_dispatch_synchronize(tk_cuda_dispatcher_t* d) {
    if (!d) return TK_ERROR_INVALID_ARGUMENT;
    DeviceScope scope(d->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    TK_CUDA_TRY(cudaStreamSynchronize(d->upload), TK_ERROR_GPU_SYNCHRONIZATION_FAILED);
    TK_CUDA_TRY(cudaStreamSynchronize(d->compute), TK_ERROR_GPU_SYNCHRONIZATION_FAILED);
    TK_CUDA_TRY(cudaStreamSynchronize(d->download), TK_ERROR_GPU_SYNCHRONIZATION_FAILED);
    return TK_SUCCESS;
}

extern "C" tk_error_code_t tk_cuda_dispatch_preprocess_image(tk_cuda_dispatcher_t* d, const tk_preprocess_params_t* p) {
    if (!d || !p) return TK_ERROR_INVALID_ARGUMENT;
    DeviceScope scope(d->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    return tk_kernels_preprocess_image(p, d->compute);
}

extern "C" tk_error_code_t tk_cuda_dispatch_depth_to_point_cloud(tk_cuda_dispatcher_t* d, const tk_depth_to_points_params_t* p) {
    if (!d || !p) return TK_ERROR_INVALID_ARGUMENT;
    DeviceScope scope(d->device_id);
    if (!scope.ok) return TK_ERROR_GPU_DEVICE_SET_FAILED;
    return tk_kernels_depth_to_point_cloud(p, d->compute);
}

// Additive accessors (no reference equivalent): raw pointer of a buffer handle and
// the dispatcher's compute stream, so tk_kernels_* launchers can be mixed in.
extern "C" void* tk_b200_buffer_device_ptr(tk_gpu_buffer_t b) { return b ? b->d_ptr : nullptr; }
extern "C" size_t tk_b200_buffer_size(tk_gpu_buffer_t b) { return b ? b->size_bytes : 0; }
extern "C" cudaStream_t tk_b200_dispatch_stream(tk_cuda_dispatcher_t* d) { return d ? d->compute : nullptr; }
extern "C" int tk_b200_dispatch_device(tk_cuda_dispatcher_t* d) { return d ? d->device_id : -1; }
