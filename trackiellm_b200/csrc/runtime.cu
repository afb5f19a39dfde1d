// runtime.cu — error reporting (src/utils/tk_error_handling.h:160-201 contract),
// device property cache and library self-description.
#include "tk_common.cuh"

#include <stdarg.h>
#include <string.h>
#include <atomic>
#include <mutex>

namespace tkb {

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
unsigned long long launch_count() { return g_launches.load(std::memory_order_relaxed); }

static std::atomic<unsigned long long*> g_trace_buf{nullptr};
static std::atomic<uint32_t> g_trace_cap{0};
TraceBuf trace_target() { return TraceBuf{g_trace_buf.load(std::memory_order_relaxed), g_trace_cap.load(std::memory_order_relaxed)}; }

bool& pdl_chain_hint() { static thread_local bool on = false; return on; }

static DeviceInfo g_info[64];
static int g_info_state[64];   // 0 unknown, 1 ok, -1 bad
static std::mutex g_info_mu;

const DeviceInfo* device_info(int* out_device) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
        cudaGetLastError();
        tk_error_set_detail("no usable CUDA device (cudaGetDevice failed)");
        return nullptr;
    }
    if (out_device) *out_device = dev;
    std::lock_guard<std::mutex> lk(g_info_mu);
    if (g_info_state[dev] == 0) {
        DeviceInfo d{};
        bool ok = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess &&
                  cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) == cudaSuccess &&
                  cudaDeviceGetAttribute(&d.cc_major, cudaDevAttrComputeCapabilityMajor, dev) == cudaSuccess;
        if (ok && d.cc_major != 10) {
            tk_error_set_detail("device %d is compute capability %d.x; libtk_b200 carries sm_100a code only", dev, d.cc_major);
            ok = false;
        }
        g_info[dev] = d;
        g_info_state[dev] = ok ? 1 : -1;
    }
    return g_info_state[dev] == 1 ? &g_info[dev] : nullptr;
}

}  // namespace tkb

static thread_local char g_detail[512];

extern "C" void tk_error_set_detail(const char* fmt, ...) {
    if (!fmt) { g_detail[0] = 0; return; }
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_detail, sizeof(g_detail), fmt, ap);
    va_end(ap);
}

extern "C" const char* tk_error_get_detail(void) { return g_detail; }

extern "C" const char* tk_error_to_string(tk_error_code_t code) {
    switch (code) {
        case TK_SUCCESS: return "success";
        case TK_ERROR_UNKNOWN: return "unknown error";
        case TK_ERROR_INVALID_ARGUMENT: return "invalid argument";
        case TK_ERROR_INVALID_STATE: return "invalid state";
        case TK_ERROR_NOT_IMPLEMENTED: return "not implemented";
        case TK_ERROR_BUFFER_TOO_SMALL: return "buffer too small";
        case TK_ERROR_NOT_INITIALIZED: return "not initialized";
        case TK_ERROR_OUT_OF_MEMORY: return "out of host memory";
        case TK_ERROR_INFERENCE_FAILED: return "inference failed";
        case TK_ERROR_INVALID_INPUT_TENSOR: return "invalid input tensor";
        case TK_ERROR_INVALID_OUTPUT_TENSOR: return "invalid output tensor";
        case TK_ERROR_GPU_ERROR: return "GPU error";
        case TK_ERROR_GPU_DEVICE_NOT_FOUND: return "no compatible GPU device";
        case TK_ERROR_GPU_CUDA_ERROR: return "CUDA runtime error";
        case TK_ERROR_GPU_KERNEL_LAUNCH: return "GPU kernel launch failed";
        case TK_ERROR_GPU_MEMORY: return "GPU memory error";
        case TK_ERROR_INVALID_DIMENSIONS: return "invalid dimensions";
        case TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE: return "CUDA kernel launch failure";
        case TK_ERROR_CUDA_KERNEL_FAILURE: return "CUDA kernel failure";
        case TK_ERROR_INVALID_CAMERA_PARAMETERS: return "invalid camera parameters";
        case TK_ERROR_GPU_DEVICE_ENUMERATION_FAILED: return "GPU device enumeration failed";
        case TK_ERROR_GPU_DEVICE_QUERY_FAILED: return "GPU device query failed";
        case TK_ERROR_GPU_DEVICE_UNSUPPORTED: return "GPU device unsupported";
        case TK_ERROR_GPU_MEMORY_INSUFFICIENT: return "insufficient GPU memory";
        case TK_ERROR_GPU_STREAM_CREATION_FAILED: return "GPU stream creation failed";
        case TK_ERROR_GPU_DEVICE_SET_FAILED: return "GPU device selection failed";
        case TK_ERROR_GPU_NOT_INITIALIZED: return "GPU dispatcher not initialized";
        case TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED: return "GPU memory allocation failed";
        case TK_ERROR_GPU_INVALID_BUFFER: return "invalid GPU buffer";
        case TK_ERROR_GPU_BUFFER_SIZE_MISMATCH: return "GPU buffer size mismatch";
        case TK_ERROR_GPU_MEMORY_COPY_FAILED: return "GPU memory copy failed";
        case TK_ERROR_GPU_SYNCHRONIZATION_FAILED: return "GPU synchronization failed";
        default: return "error";
    }
}

extern "C" const char* tk_b200_version(void) { return "tk-b200-vision 0.1 (sm_100a)"; }

// L2 fetch granularity hint of the current device's context (bytes: 32, 64 or 128;
// 0 = query only).  Returns the value in effect, or -1 on error.  Sector-strided
// scans (decode pass 1 reads 4 bytes of every 340-byte row) move fewer DRAM bytes
// with a small granularity; streaming kernels are indifferent.
extern "C" int tk_b200_l2_fetch_granularity(int bytes) {
    if (bytes > 0 && cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes) != cudaSuccess) { cudaGetLastError(); return -1; }
    size_t v = 0;
    if (cudaDeviceGetLimit(&v, cudaLimitMaxL2FetchGranularity) != cudaSuccess) { cudaGetLastError(); return -1; }
    return (int)v;
}

// Debug trace target: a DEVICE buffer of `bytes` (>= 128) that the caller zeroes and reads back; NULL / 0 turns it off.
extern "C" void tk_b200_debug_trace(void* d_buf, size_t bytes) {
    const uint32_t cap = (d_buf && bytes >= 128) ? (uint32_t)((bytes - 64) / 64) : 0u;
    tkb::g_trace_cap.store(cap, std::memory_order_relaxed);
    tkb::g_trace_buf.store(cap ? static_cast<unsigned long long*>(d_buf) : nullptr, std::memory_order_relaxed);
}

extern "C" unsigned long long tk_b200_launch_count(void) { return tkb::launch_count(); }

extern "C" int tk_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}
