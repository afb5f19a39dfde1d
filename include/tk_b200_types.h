/*
 * tk_b200_types.h — the C types the vision hot path sits behind.
 *
 * Every type here is a layout-compatible restatement (same field order, same
 * field types, same enumerator values) of a type the reference declares, so a
 * reference translation unit can be relinked against libtk_b200.so unchanged.
 * Where the reference's .cu sources use enumerators its enum never declares,
 * they are added at the END of the 5000 band so every existing value keeps its
 * number.  Citations are relative to the reference tree.
 */
#ifndef TK_B200_TYPES_H
#define TK_B200_TYPES_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__) || defined(__clang__)
#define TK_NODISCARD __attribute__((warn_unused_result))
#else
#define TK_NODISCARD
#endif

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes: src/utils/tk_error_handling.h:40-124 -------------------- */
typedef enum tk_error_code_t {
    TK_SUCCESS = 0,

    TK_ERROR_UNKNOWN = 1000,
    TK_ERROR_INVALID_ARGUMENT,
    TK_ERROR_INVALID_STATE,
    TK_ERROR_NOT_IMPLEMENTED,
    TK_ERROR_BUFFER_TOO_SMALL,
    TK_ERROR_TIMEOUT,
    TK_ERROR_PERMISSION_DENIED,
    TK_ERROR_NOT_INITIALIZED,

    TK_ERROR_OUT_OF_MEMORY = 2000,
    TK_ERROR_MEMORY_ALIGNMENT,
    TK_ERROR_MEMORY_POOL_EXHAUSTED,
    TK_ERROR_MEMORY_DOUBLE_FREE,
    TK_ERROR_MEMORY_INVALID_POINTER,

    TK_ERROR_IO = 3000,
    TK_ERROR_FILE_NOT_FOUND,
    TK_ERROR_FILE_READ,
    TK_ERROR_FILE_WRITE,
    TK_ERROR_FILE_CORRUPT,
    TK_ERROR_CONFIG_PARSE_FAILED,

    TK_ERROR_MODEL_LOAD_FAILED = 4000,
    TK_ERROR_MODEL_VERIFICATION_FAILED,
    TK_ERROR_INFERENCE_FAILED,
    TK_ERROR_INVALID_INPUT_TENSOR,
    TK_ERROR_INVALID_OUTPUT_TENSOR,
    TK_ERROR_BACKEND_NOT_SUPPORTED,

    TK_ERROR_GPU_ERROR = 5000,
    TK_ERROR_GPU_DEVICE_NOT_FOUND,
    TK_ERROR_GPU_DRIVER_VERSION,
    TK_ERROR_GPU_CUDA_ERROR,
    TK_ERROR_GPU_METAL_ERROR,
    TK_ERROR_GPU_ROCM_ERROR,
    TK_ERROR_GPU_KERNEL_LAUNCH,
    TK_ERROR_GPU_MEMORY,
    /* Additive: names the reference's CUDA sources use but never declare
     * (src/gpu/cuda/tk_cuda_kernels.cu:38,273,299,471 and
     *  src/gpu/cuda/tk_cuda_dispatch.cu:75-477). */
    TK_ERROR_INVALID_DIMENSIONS,
    TK_ERROR_CUDA_KERNEL_LAUNCH_FAILURE,
    TK_ERROR_CUDA_KERNEL_FAILURE,
    TK_ERROR_INVALID_CAMERA_PARAMETERS,
    TK_ERROR_GPU_DEVICE_ENUMERATION_FAILED,
    TK_ERROR_GPU_DEVICE_QUERY_FAILED,
    TK_ERROR_GPU_DEVICE_UNSUPPORTED,
    TK_ERROR_GPU_MEMORY_INSUFFICIENT,
    TK_ERROR_GPU_STREAM_CREATION_FAILED,
    TK_ERROR_GPU_DEVICE_SET_FAILED,
    TK_ERROR_GPU_NOT_INITIALIZED,
    TK_ERROR_GPU_MEMORY_ALLOCATION_FAILED,
    TK_ERROR_GPU_INVALID_BUFFER,
    TK_ERROR_GPU_BUFFER_SIZE_MISMATCH,
    TK_ERROR_GPU_MEMORY_COPY_FAILED,
    TK_ERROR_GPU_SYNCHRONIZATION_FAILED,

    TK_ERROR_NETWORK_ERROR = 6000,
    TK_ERROR_CONNECTION_FAILED,
    TK_ERROR_CONNECTION_CLOSED,
    TK_ERROR_DNS_RESOLUTION_FAILED,
    TK_ERROR_SOCKET_ERROR,

    TK_ERROR_THREAD_CREATE_FAILED = 7000,
    TK_ERROR_MUTEX_ERROR,
    TK_ERROR_SEMAPHORE_ERROR,
    TK_ERROR_TASK_QUEUE_FULL,
    TK_ERROR_FUTURE_CANCELLED,

    TK_ERROR_FFI_PANIC = 8000,
    TK_ERROR_FFI_INVALID_STRING,

    TK_ERROR_CODE_COUNT
} tk_error_code_t;

/* src/utils/tk_error_handling.h:160,185,201 */
const char* tk_error_to_string(tk_error_code_t code);
void        tk_error_set_detail(const char* fmt, ...);
const char* tk_error_get_detail(void);

/* ---- frame descriptor: src/cortex/tk_cortex_main.h:98-104 ----------------- */
typedef enum {
    TK_PIXEL_FORMAT_RGB8  = 0,
    TK_PIXEL_FORMAT_RGBA8 = 1,
    /* Additive: planar I420 container of src/ffi/c_api/tk_ffi_api.c:917-931
     * (Y w*h, then U, then V, (w/2)*(h/2) each, one allocation). */
    TK_PIXEL_FORMAT_I420  = 2,
    /* Additive: packed B,G,R byte order (TK_VISION_FMT_BGR24 of the FFI frame
     * container, src/ffi/c_api/tk_ffi_api.h).  The tensor planes stay R, G, B. */
    TK_PIXEL_FORMAT_BGR8  = 3
} tk_pixel_format_e;

typedef struct tk_video_frame_s {
    uint32_t width;
    uint32_t height;
    uint32_t stride;            /* bytes per row (I420: luma row pitch)            */
    tk_pixel_format_e format;   /* the reference uses an anonymous enum, same ABI  */
    const uint8_t* data;
} tk_video_frame_t;

/* ---- src/vision/tk_vision_pipeline.h:158-163 ------------------------------ */
typedef struct {
    int x;
    int y;
    int w;
    int h;
} tk_rect_t;

/* ---- src/vision/tk_object_detector.h:66-71 -------------------------------- */
typedef struct {
    uint32_t    class_id;
    const char* label;       /* points into the caller's class_labels table */
    float       confidence;
    tk_rect_t   bbox;
} tk_detection_result_t;

/* ---- src/vision/tk_vision_pipeline.h:56-61 -------------------------------- */
typedef enum {
    TK_VISION_BACKEND_CPU,
    TK_VISION_BACKEND_CUDA,
    TK_VISION_BACKEND_METAL,
    TK_VISION_BACKEND_ROCM
} tk_vision_backend_e;

/* ---- src/vision/tk_object_detector.h:35-58 -------------------------------- */
typedef struct tk_path_s tk_path_t;                       /* src/internal_tools/tk_file_manager.h:40 (opaque) */
typedef struct tk_object_detector_s tk_object_detector_t;
typedef struct {
    tk_vision_backend_e backend;
    int                 gpu_device_id;
    tk_path_t*          model_path;
    uint32_t            input_width;
    uint32_t            input_height;
    const char**        class_labels;   /* not copied; must outlive the detector */
    size_t              class_count;
    float               confidence_threshold;
    float               iou_threshold;
} tk_object_detector_config_t;

/* ---- src/gpu/tk_gpu_helper.h:25-27,42-57,67-80,91-105,119-130 ------------- */
typedef struct {
    float x, y, z;
} TkFloat3;

typedef struct {
    const void* d_input_image;
    uint32_t input_width;
    uint32_t input_height;
    uint32_t input_stride_bytes;
    void* d_output_tensor;
    uint32_t output_width;
    uint32_t output_height;
    TkFloat3 mean;
    TkFloat3 std_dev;
} tk_preprocess_params_t;

typedef struct {
    const void* d_raw_depth_map;
    uint32_t width;
    uint32_t height;
    void* d_metric_depth_map;
    float scale;
    float shift;
} tk_postprocess_depth_params_t;

typedef struct {
    const void* d_metric_depth_map;
    uint32_t width;
    uint32_t height;
    void* d_point_cloud;
    float fx;
    float fy;
    float cx;
    float cy;
} tk_depth_to_points_params_t;

typedef struct {
    const void* d_input_tensor;
    void* d_output_tensor;
    uint32_t num_rows;
    uint32_t num_cols;
} tk_softmax_params_t;

/* ---- opaque handles: src/gpu/cuda/tk_cuda_dispatch.h:35-49 ---------------- */
typedef struct tk_cuda_dispatcher_s tk_cuda_dispatcher_t;
typedef struct tk_gpu_buffer_s* tk_gpu_buffer_t;
typedef struct tk_gpu_event_s*  tk_gpu_event_t;

typedef struct {
    int device_id;
} tk_cuda_dispatcher_config_t;

/* ---- src/gpu/extensions/cuda/tk_cuda_tensor_ops.h:54-86 ------------------- */
typedef enum {
    TK_TENSOR_TYPE_F32 = 0,
    TK_TENSOR_TYPE_F16,
    TK_TENSOR_TYPE_I32,
    TK_TENSOR_TYPE_I8,
    TK_TENSOR_TYPE_U8
} tk_tensor_data_type_t;

typedef enum {
    TK_TENSOR_LAYOUT_NCHW = 0,
    TK_TENSOR_LAYOUT_NHWC,
    TK_TENSOR_LAYOUT_NC,
    TK_TENSOR_LAYOUT_NCW
} tk_tensor_layout_t;

typedef struct {
    tk_gpu_buffer_t buffer;
    tk_tensor_data_type_t type;
    tk_tensor_layout_t layout;
    uint32_t dimensions;
    uint32_t shape[8];
    size_t stride[8];
    size_t data_size_bytes;
    uint32_t batch_size;
} tk_tensor_descriptor_t;

/* CUDA stream handle without pulling in cuda_runtime.h for C callers. */
#if !defined(__CUDA_RUNTIME_H__) && !defined(__DRIVER_TYPES_H__)
typedef struct CUstream_st* cudaStream_t;
#endif

#ifdef __cplusplus
}
#endif
#endif /* TK_B200_TYPES_H */
