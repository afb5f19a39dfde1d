/*
 * tk_b200_vision.h — C-ABI of libtk_b200.so, the B200 (sm_100a) implementation
 * of TrackieLLM's vision hot path.
 *
 * Part 1 keeps the reference's own entry points (name, signature, argument
 * meaning, error behaviour); each cites the reference declaration it replaces.
 * Part 2 is additive (batched / fused forms the reference has no equivalent
 * of), same conventions: every call returns tk_error_code_t, TK_SUCCESS == 0,
 * never aborts; kernel launchers are asynchronous on the given stream.
 *
 * There is NO CPU fallback behind any of these: without a CUDA device every
 * compute entry point returns a TK_ERROR_GPU_* code.
 */
#ifndef TK_B200_VISION_H
#define TK_B200_VISION_H

#include "tk_b200_types.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ======================================================================== */
/* Part 1 — reference entry points                                           */
/* ======================================================================== */

/* src/vision/tk_image_preprocessor.h:49-56.  HOST pointers in and out; the
 * frame is uploaded, resized on the GPU with the reference CPU mapping
 * (src/vision/tk_image_preprocessor.c:43-69), and the CHW fp32 tensor is
 * downloaded; bit-identical to the reference's AVX build.  Unlike the
 * reference it honours frame->stride (0 means tightly packed) and rejects
 * width/height < 2 (the reference reads out of bounds there). */
TK_NODISCARD tk_error_code_t tk_preprocessor_resize_and_normalize_to_chw(
    const tk_video_frame_t* frame, float* out_tensor, uint32_t target_width,
    uint32_t target_height, const float mean[3], const float std_dev[3]);

/* src/gpu/cuda/tk_cuda_kernels.h:51.  DEVICE pointers inside params.  Uses the
 * reference CPU mapping, not the half-pixel mapping of the reference CUDA
 * kernel (src/gpu/cuda/tk_cuda_kernels.cu:133-158): the north star pins the
 * CPU implementation.  Output fp32 CHW. */
TK_NODISCARD tk_error_code_t tk_kernels_preprocess_image(const tk_preprocess_params_t* params, cudaStream_t stream);

/* src/gpu/cuda/tk_cuda_kernels.h:64,78 */
TK_NODISCARD tk_error_code_t tk_kernels_postprocess_depth_map(const tk_postprocess_depth_params_t* params, cudaStream_t stream);
TK_NODISCARD tk_error_code_t tk_kernels_depth_to_point_cloud(const tk_depth_to_points_params_t* params, cudaStream_t stream);

/* src/gpu/cuda/tk_cuda_kernels.h:92.  Any num_cols (the reference rejects
 * > 1024 and is wrong for non-power-of-two, tk_cuda_kernels.cu:331-336,369). */
TK_NODISCARD tk_error_code_t tk_kernels_softmax(const tk_softmax_params_t* params, cudaStream_t stream);

/* src/gpu/cuda/tk_cuda_dispatch.h:72-171 */
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_create(tk_cuda_dispatcher_t** out_dispatcher, const tk_cuda_dispatcher_config_t* config);
void tk_cuda_dispatch_destroy(tk_cuda_dispatcher_t** dispatcher);
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_malloc(tk_cuda_dispatcher_t* dispatcher, tk_gpu_buffer_t* out_buffer, size_t size_bytes);
void tk_cuda_dispatch_free(tk_cuda_dispatcher_t* dispatcher, tk_gpu_buffer_t* buffer);
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_upload_async(tk_cuda_dispatcher_t* dispatcher, tk_gpu_buffer_t dst_buffer, const void* src_host_ptr, size_t size_bytes);
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_download_async(tk_cuda_dispatcher_t* dispatcher, void* dst_host_ptr, tk_gpu_buffer_t src_buffer, size_t size_bytes);
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_synchronize(tk_cuda_dispatcher_t* dispatcher);
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_preprocess_image(tk_cuda_dispatcher_t* dispatcher, const tk_preprocess_params_t* params);
TK_NODISCARD tk_error_code_t tk_cuda_dispatch_depth_to_point_cloud(tk_cuda_dispatcher_t* dispatcher, const tk_depth_to_points_params_t* params);

/* src/gpu/extensions/cuda/tk_cuda_tensor_ops.h:513-522.  boxes: f32 [.., K, 4]
 * as (x, y, w, h) top-left form, truncated to integer rects like
 * src/vision/tk_object_detector.c:347-350; scores: f32 [K]; class-agnostic
 * (no class tensor in this signature); indices: device i32[max_detections];
 * num_detections: HOST pointer (the call synchronises). */
TK_NODISCARD tk_error_code_t tk_cuda_tensor_nms(
    tk_cuda_dispatcher_t* dispatcher, const tk_tensor_descriptor_t* boxes,
    const tk_tensor_descriptor_t* scores, tk_gpu_buffer_t indices,
    uint32_t* num_detections, uint32_t max_detections, float iou_threshold,
    float score_threshold);

/* src/vision/tk_attribute_classifier.h (tk_classify_dominant_color,
 * tk_classify_door_state).  HOST frame; out string is malloc'd, caller frees. */
TK_NODISCARD tk_error_code_t tk_classify_dominant_color(const tk_video_frame_t* frame, const tk_rect_t* bbox, char** out_color_name);
TK_NODISCARD tk_error_code_t tk_classify_door_state(const tk_video_frame_t* frame, const tk_rect_t* bbox, char** out_state_name);

/* src/vision/tk_object_detector.h:96,102,138-143,157,166.  Same signatures and
 * error behaviour.  The reference's create() opens an ONNX Runtime session on
 * config->model_path (src/vision/tk_object_detector.c:93-120); ONNX Runtime is a
 * third-party engine outside this path, so here the network is BOUND to the
 * detector with tk_b200_object_detector_bind_inference (Part 2) and receives
 * DEVICE pointers — the I/O-binding form of the reference's host-tensor call
 * (:266-294) — so pre-process -> network -> post-process never leaves HBM.
 * backend must be TK_VISION_BACKEND_CUDA (no CPU fallback: TK_ERROR_NOT_IMPLEMENTED).
 * detect(): upload, fused pre-process (fp32, ImageNet mean/std of :248-249),
 * bound network, decode + class-wise NMS, boxes restored to frame px, results
 * malloc'd for tk_object_detector_free_results.  Not thread-safe per detector
 * (tk_object_detector.h:134-137). */
TK_NODISCARD tk_error_code_t tk_object_detector_create(tk_object_detector_t** out_detector, const tk_object_detector_config_t* config);
void tk_object_detector_destroy(tk_object_detector_t** detector);
TK_NODISCARD tk_error_code_t tk_object_detector_detect(tk_object_detector_t* detector, const tk_video_frame_t* video_frame,
                                                       tk_detection_result_t** out_results, size_t* out_result_count);
void tk_object_detector_free_results(tk_detection_result_t** results);
void tk_object_detector_update_thresholds(tk_object_detector_t* detector, float confidence_threshold, float iou_threshold);

/* ======================================================================== */
/* Part 2 — additive entry points                                            */
/* ======================================================================== */

typedef enum { TK_B200_OUT_F32 = 0, TK_B200_OUT_F16 = 1 } tk_b200_out_type_e;
typedef enum { TK_B200_RESIZE_STRETCH = 0, TK_B200_RESIZE_LETTERBOX = 1 } tk_b200_resize_mode_e;
/* EXACT: every fp32 operation of the reference in its order, no contraction;
 *        bit-identical to the reference's AVX build.
 * FAST : lerp form with FMA; within 1e-5 + 1e-5|ref| of the reference.
 * HALF : fp16 output only (RGB8 / RGBA8 input; otherwise FAST is used): bytes
 *        become exact half2 pairs centred on 128 and the lerps run in HFMA2;
 *        within HALF of the 1e-2 + 1e-2|ref| fp16 bar of the reference for any
 *        content (bound derived in DESIGN.md §3, asserted on adversarial frames).
 * EXACT is what every handle and launcher uses unless told otherwise. */
typedef enum { TK_B200_ARITH_EXACT = 0, TK_B200_ARITH_FAST = 1, TK_B200_ARITH_HALF = 2 } tk_b200_arith_e;
typedef enum { TK_B200_ACT_RAW = 0, TK_B200_ACT_SIGMOID = 1, TK_B200_ACT_SOFTMAX = 2 } tk_b200_activation_e;

/* Batched fused pre-process: N frames of one geometry -> [N][3][H][W]. */
typedef struct {
    const void*        d_input;                 /* frame 0 (device)                                  */
    const void* const* d_input_ptrs;            /* optional DEVICE table of N frame pointers         */
    uint64_t           input_frame_pitch_bytes; /* distance between frames when d_input_ptrs == NULL */
    uint32_t           num_frames;
    uint32_t           input_width;
    uint32_t           input_height;
    uint32_t           input_stride_bytes;      /* 0 = tightly packed                                */
    tk_pixel_format_e  input_format;            /* RGB8, RGBA8 or I420                               */
    void*              d_output;                /* device, fp32 or fp16, NCHW                        */
    uint32_t           output_width;
    uint32_t           output_height;
    tk_b200_out_type_e output_type;
    tk_b200_resize_mode_e resize_mode;
    float              pad_value;               /* letterbox fill on the 0..255 scale                */
    tk_b200_arith_e    arith;
    TkFloat3           mean;
    TkFloat3           std_dev;
} tk_b200_preprocess_batch_params_t;

/* Asynchronous on `stream`.  Large launches hand part of their output rows out dynamically from a chunk counter in the
 * library's device memory; every call takes the next of 64 counters, so up to 64 launches may be in flight on a device
 * at once.  A launch recorded into a CUDA graph keeps the counter it was captured with: replays of ONE graph must not
 * overlap each other on the GPU (replay it on one stream, or capture it once per stream). */
TK_NODISCARD tk_error_code_t tk_b200_preprocess_batch(const tk_b200_preprocess_batch_params_t* params, cudaStream_t stream);

/* One decoded proposal / detection (28 bytes, packed ints and one float). */
typedef struct {
    int32_t   index;       /* proposal row inside its frame */
    tk_rect_t bbox;
    float     confidence;
    uint32_t  class_id;
} tk_b200_candidate_t;

/* YOLO-head decode + threshold + ordered compaction
 * (src/vision/tk_object_detector.c:343-355 completed). */
typedef struct {
    const float*         d_logits;          /* [N][P][5+C] fp32 row-major                         */
    uint32_t             num_frames;
    uint32_t             num_proposals;
    uint32_t             num_classes;
    float                confidence_threshold;
    tk_b200_activation_e activation;
    tk_b200_candidate_t* d_candidates;      /* [N][max_candidates], ascending proposal index      */
    uint32_t*            d_counts;          /* [N] rows passing the threshold (may exceed the cap) */
    uint32_t             max_candidates;
} tk_b200_decode_params_t;

TK_NODISCARD tk_error_code_t tk_kernels_decode_detections(const tk_b200_decode_params_t* params, cudaStream_t stream);

/* Class-wise greedy IoU NMS over per-frame candidate lists. */
typedef struct {
    const tk_b200_candidate_t* d_candidates;   /* [N][max_candidates]                       */
    const uint32_t*            d_counts;       /* [N]; min(count, max_candidates) are used   */
    uint32_t                   num_frames;
    uint32_t                   max_candidates;
    float                      iou_threshold;
    uint32_t                   max_detections;
    int                        class_agnostic; /* 0 = only equal class_id suppresses         */
    tk_b200_candidate_t*       d_detections;   /* [N][max_detections], (score desc, index asc) */
    uint32_t*                  d_num_detections; /* [N]                                      */
    /* Optional coordinate restore of the kept boxes, model-input px -> source-frame
     * px (completes the empty convert_to_original_coordinates,
     * src/vision/tk_object_detector.c:76-79,365,372-375): x,w scaled by
     * src_width/in_width, y,h by src_height/in_height in fp32, truncated.  All four
     * zero = boxes stay in model-input px. */
    uint32_t                   restore_src_width, restore_src_height;
    uint32_t                   restore_in_width, restore_in_height;
} tk_b200_nms_params_t;

TK_NODISCARD tk_error_code_t tk_kernels_nms_classwise(const tk_b200_nms_params_t* params, cudaStream_t stream);

/* The same NMS for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same parameters and
 * results, bit for bit, as tk_kernels_nms_classwise; calls no cudaMalloc / cudaFree / stream-ordered allocator: the
 * per-CTA work areas live in a file-scope device array of 6.1 MiB (8 x nms work area of 16384 candidates) that every
 * GPU that loads the library holds until the process exits.  max_candidates above
 * tk_b200_nms_classwise_static_max_candidates() (16384) -> TK_ERROR_BUFFER_TOO_SMALL with a detail text.
 * All calls on a device share the array: calls that overlap on the GPU (two streams, two threads) race - the caller
 * serialises them.  Nothing in the array has to be cleared between calls (every frame initialises what it reads). */
TK_NODISCARD tk_error_code_t tk_b200_nms_classwise_static(const tk_b200_nms_params_t* params, cudaStream_t stream);
uint32_t tk_b200_nms_classwise_static_max_candidates(void);

/* Decode + class-wise NMS of whole frames in ONE kernel (completes src/vision/tk_object_detector.c:343-357 in one
 * pass).  Same results, bit for bit, as tk_kernels_decode_detections followed by tk_kernels_nms_classwise;
 * d_candidates / d_counts receive the intermediate list exactly as decode leaves it.  Two layouts:
 *   0  a thread-block cluster of 8 CTAs per frame decodes it, then the cluster's first CTA runs the frame's NMS at once
 *      (one launch, no gap, the NMS of early frames hides under the decode of late ones);
 *   1  one persistent 256-thread CTA (<= 32 registers, 25 KB of shared memory) works whole frames — cut to be
 *      RESIDENT BESIDE another kernel's CTAs on a second stream (measured in DESIGN.md §9; not the default).
 * Frames with more than 512 candidates run their NMS through d_workspace (global memory): slower, same result. */
typedef struct {
    const float*         d_logits;            /* [N][P][5+C] fp32 row-major                               */
    uint32_t             num_frames, num_proposals, num_classes;
    float                confidence_threshold;
    tk_b200_activation_e activation;
    tk_b200_candidate_t* d_candidates;        /* [N][max_candidates]                                      */
    uint32_t*            d_counts;            /* [N] rows passing the threshold (may exceed the cap)      */
    uint32_t             max_candidates;      /* any (no 4096 limit: larger lists sort in d_workspace)    */
    float                iou_threshold;
    uint32_t             max_detections;
    int                  class_agnostic;
    tk_b200_candidate_t* d_detections;        /* [N][max_detections], (score desc, index asc)             */
    uint32_t*            d_num_detections;    /* [N]                                                      */
    uint32_t             restore_src_width, restore_src_height, restore_in_width, restore_in_height;  /* as tk_b200_nms_params_t */
    void*                d_workspace;         /* needed when max_candidates > 512                         */
    size_t               workspace_bytes;     /* >= tk_b200_postprocess_fused_workspace_bytes(max_candidates, num_frames, layout) */
    int                  layout;              /* 0 cluster per frame, 1 persistent co-resident CTAs       */
} tk_b200_post_params_t;

size_t tk_b200_postprocess_fused_workspace_bytes(uint32_t max_candidates, uint32_t num_frames, int layout);
TK_NODISCARD tk_error_code_t tk_kernels_postprocess_fused(const tk_b200_post_params_t* params, cudaStream_t stream);

/* Row softmax with fused argmax (either output may be NULL). */
typedef struct {
    const float* d_input;       /* [rows][cols]           */
    float*       d_softmax;     /* [rows][cols] or NULL   */
    uint32_t*    d_argmax;      /* [rows] or NULL         */
    uint32_t     num_rows;
    uint32_t     num_cols;
} tk_b200_argmax_params_t;

TK_NODISCARD tk_error_code_t tk_kernels_argmax_rows(const tk_b200_argmax_params_t* params, cudaStream_t stream);

/* Per-detection dominant colour (src/vision/tk_attribute_classifier.c:40-97)
 * for M boxes of one device-resident RGB8 frame. */
typedef struct {
    const uint8_t*   d_frame;
    uint32_t         width, height, stride_bytes;
    const tk_rect_t* d_boxes;        /* [M]                              */
    uint32_t         num_boxes;
    uint32_t*        d_color_ids;    /* [M] index into tk_b200_color_name */
    unsigned long long* d_bins;      /* [M][9] histogram or NULL          */
} tk_b200_color_params_t;

TK_NODISCARD tk_error_code_t tk_kernels_dominant_color(const tk_b200_color_params_t* params, cudaStream_t stream);
const char* tk_b200_color_name(uint32_t color_id);

/* Per-detection door state (src/vision/tk_attribute_classifier.c:100-157):
 * strong-vertical-edge density of the box interior; state 1 = "closed". */
typedef struct {
    const uint8_t*   d_frame;
    uint32_t         width, height, stride_bytes;
    const tk_rect_t* d_boxes;        /* [M]              */
    uint32_t         num_boxes;
    uint32_t*        d_states;       /* [M] 1 = closed   */
    int32_t*         d_edge_pixels;  /* [M] or NULL      */
} tk_b200_door_params_t;

TK_NODISCARD tk_error_code_t tk_kernels_door_state(const tk_b200_door_params_t* params, cudaStream_t stream);

/* MiDaS inverse depth -> metric depth for a batch of maps
 * (convert_inverse_depth_to_metric, src/vision/tk_depth_midas.c:473-499; the
 * reference passes TK_DEPTH_ESTIMATOR_MIN/MAX_DEPTH = 0.1 / 10.0, :46-47).
 * Each map is normalised by its own min/max.  In place (same pointer) is
 * allowed for maps of up to 196608 elements. */
typedef struct {
    const float* d_inverse_depth;   /* [num_maps][height][width]  */
    float*       d_metric_depth;    /* [num_maps][height][width]  */
    uint32_t     num_maps;
    uint32_t     width, height;
    float        min_depth, max_depth;
} tk_b200_depth_metric_params_t;

TK_NODISCARD tk_error_code_t tk_kernels_inverse_depth_to_metric(const tk_b200_depth_metric_params_t* params, cudaStream_t stream);

/* Per-detection distance from a metric depth map (calculate_raw_distance,
 * src/vision/src/object_analysis.rs:227-279): bbox -> depth-map window ->
 * samples with 0.1 < d < 100 -> sort -> IQR fence -> mean; -1.0 when the
 * window is degenerate or holds fewer than 10 valid samples.  Bit-identical
 * to the sequential fp32 reference.  Frame f uses depth map f. */
typedef struct {
    const tk_b200_candidate_t* d_detections;     /* [N][max_detections] (bbox in frame px)   */
    const uint32_t*            d_num_detections; /* [N] or NULL = max_detections everywhere  */
    uint32_t                   num_frames;       /* N                                        */
    uint32_t                   max_detections;
    const float*               d_depth_maps;     /* [N][depth_height][depth_width]           */
    uint32_t                   depth_width, depth_height;
    uint32_t                   frame_width, frame_height;   /* px space of the boxes        */
    float*                     d_distances;      /* [N][max_detections]; unused slots = -1   */
    void*                      d_workspace;      /* needed when depth_width*depth_height > 16384 */
    size_t                     workspace_bytes;  /* >= tk_b200_detection_distance_workspace_bytes */
} tk_b200_distance_params_t;

size_t tk_b200_detection_distance_workspace_bytes(uint32_t depth_width, uint32_t depth_height);
TK_NODISCARD tk_error_code_t tk_kernels_detection_distance(const tk_b200_distance_params_t* params, cudaStream_t stream);

/* ------------------------------------------------------------------------ */
/* Batched pipeline: pre-process + decode + class-wise NMS for a batch of     */
/* frames of one geometry (the step BASELINE.json's metric counts).  The      */
/* reference runs this one frame per call inside tk_object_detector_detect    */
/* (src/vision/tk_object_detector.c:182-203: preprocess_frame :223-259,       */
/* postprocess_detections :303-368); thresholds as tk_object_detector_config_t */
/* (src/vision/tk_object_detector.h:44-58).                                    */
/* ------------------------------------------------------------------------ */
typedef struct tk_b200_pipeline_s tk_b200_pipeline_t;

typedef struct {
    int                   device_id;
    uint32_t              max_batch;          /* most frames one process call may carry         */
    uint32_t              chunk_frames;       /* frames per in-flight chunk of process_host; 0 = 8 */
    /* frame geometry (all frames of a pipeline share it) */
    uint32_t              input_width, input_height;
    uint32_t              input_stride_bytes; /* 0 = tightly packed                              */
    tk_pixel_format_e     input_format;       /* RGB8, RGBA8, I420                               */
    /* model input tensor */
    uint32_t              tensor_width, tensor_height;   /* 640 x 640, tk_vision_pipeline.c:368-378 */
    tk_b200_out_type_e    tensor_type;
    tk_b200_resize_mode_e resize_mode;
    tk_b200_arith_e       arith;
    float                 pad_value;
    TkFloat3              mean, std_dev;      /* tk_object_detector.c:248-249                    */
    /* detector head */
    uint32_t              num_proposals, num_classes;    /* 25200, 80                             */
    tk_b200_activation_e  activation;
    float                 confidence_threshold;          /* tk_object_detector.h:55               */
    float                 iou_threshold;                 /* tk_object_detector.h:56               */
    uint32_t              max_candidates;     /* pre-NMS cap per frame (<= 4096); 0 = 4096       */
    uint32_t              max_detections;     /* post-NMS cap per frame; 0 = 500 (tk_object_detector.c:39) */
    int                   class_agnostic;
    int                   restore_coordinates; /* 1 = boxes returned in source-frame px          */
    int                   stage_host_logits;  /* process_host with page-locked host logits (pageable logits are always
                                                 staged): 0 = the rows stay in host memory — their objectness column
                                                 is packed by host threads and copied as one dense array, decode reads
                                                 only the surviving rows over PCIe (batches under 65536 rows behave as 2); 1 = copy the whole tensor to HBM
                                                 first; 2 = decode reads objectness words and surviving rows in place
                                                 over PCIe (one read request per row)                              */
    int                   fuse_postprocess;   /* 0 / 1 = decode_kernel, then nms_kernel, behind the pre-process kernel (the
                                                 default: fastest as measured); decode + NMS as ONE kernel
                                                 (tk_kernels_postprocess_fused): 3 = a cluster of 8 CTAs per frame decodes it
                                                 and then runs its NMS, 2 = the persistent co-resident form on a second stream
                                                 BESIDE the pre-process kernel.  Both measured slower, kept
                                                 selectable; results are identical in every mode.                          */
} tk_b200_pipeline_config_t;

TK_NODISCARD tk_error_code_t tk_b200_pipeline_create(tk_b200_pipeline_t** out_pipeline, const tk_b200_pipeline_config_t* config);
void tk_b200_pipeline_destroy(tk_b200_pipeline_t** pipeline);
/* src/vision/tk_object_detector.h:166 (tk_object_detector_update_thresholds) */
TK_NODISCARD tk_error_code_t tk_b200_pipeline_update_thresholds(tk_b200_pipeline_t* pipeline, float confidence_threshold, float iou_threshold);

/* Candidate-cap report.  max_candidates is a cap the CALLER chooses (the reference keeps every row over the threshold,
 * src/vision/tk_object_detector.c:339-355; the reference-typed entry points tk_object_detector_detect and
 * tk_b200_postprocess_detections therefore never truncate: a frame over the kernels' 4096 is done again uncapped).
 * A frame of a pipeline call with more rows over the threshold than max_candidates is cut to the first
 * max_candidates rows by proposal index before NMS; these two calls say when that happened.
 *   _candidate_overflow: after process_host[_frames] — frames of that call that were cut, largest row count seen
 *   _candidate_counts_device_ptr: process_device — u32[n] rows over the threshold per frame (device, stream-ordered) */
TK_NODISCARD tk_error_code_t tk_b200_pipeline_candidate_overflow(tk_b200_pipeline_t* pipeline, uint32_t* out_frames, uint32_t* out_max_count);
const uint32_t* tk_b200_pipeline_candidate_counts_device_ptr(tk_b200_pipeline_t* pipeline);

/* DEVICE-resident inputs and outputs; asynchronous on `stream` (3 launches; 2 with a fused post-process).
 * d_frames [n][frame bytes], d_logits [n][P][5+C] f32, d_tensor [n][3][H][W],
 * d_detections [n][max_detections], d_num_detections [n]. */
TK_NODISCARD tk_error_code_t tk_b200_pipeline_process_device(
    tk_b200_pipeline_t* pipeline, const void* d_frames, const float* d_logits, uint32_t num_frames,
    void* d_tensor, tk_b200_candidate_t* d_detections, uint32_t* d_num_detections, cudaStream_t stream);

/* HOST inputs and outputs (pinned memory from tk_b200_host_alloc copies by DMA;
 * pageable memory works, slower).  The batch is cut into chunks; upload of chunk
 * k+1, compute of chunk k and download of chunk k-1 overlap on three streams
 * ordered by events.  Page-locked h_logits are not copied: the 4-byte objectness
 * word of each 340-byte row is packed into a dense column on the host (threads,
 * while the frame copies run) and uploaded, and decode reads only the surviving
 * rows in place over PCIe — 2 % of the tensor crosses the link.  Page-locked
 * h_frames whose geometry only samples two source rows per output row
 * (height ratio >= 2, e.g. 4K -> 640) and whose rows are multiples of 16 bytes
 * are not copied either: the pre-process kernel's row ring fetches exactly
 * those rows in place (4K I420: 8.6 of 12.4 MB per frame cross the link).  h_logits may
 * be NULL (pre-process only) and h_frames may be NULL (post-process only).  The model-input tensors stay on the device for
 * the detector (tk_b200_pipeline_tensor_device_ptr) and are also downloaded when
 * h_tensor != NULL.  Blocks until the outputs are in host memory. */
TK_NODISCARD tk_error_code_t tk_b200_pipeline_process_host(
    tk_b200_pipeline_t* pipeline, const void* h_frames, const float* h_logits, uint32_t num_frames,
    void* h_tensor, tk_b200_candidate_t* h_detections, uint32_t* h_num_detections);

/* As process_host, but the logits are already on the device (where a co-located
 * detector leaves them).  d_logits must be COMPLETE and visible to the
 * device before the call (synchronise the stream that wrote them, or record an event and
 * wait for it): the call reads them on the pipeline's own streams. */
TK_NODISCARD tk_error_code_t tk_b200_pipeline_process_host_frames(
    tk_b200_pipeline_t* pipeline, const void* h_frames, const float* d_logits, uint32_t num_frames,
    void* h_tensor, tk_b200_candidate_t* h_detections, uint32_t* h_num_detections);

/* Per-stage timing of process_device with CUDA events recorded on the caller's
 * stream around each of the three launches: stage_timing(p, max_calls) arms it
 * for the next max_calls calls (0 disarms); stage_times sums the recorded
 * durations in ms — [0] pre-process, [1] decode, [2] NMS. */
TK_NODISCARD tk_error_code_t tk_b200_pipeline_stage_timing(tk_b200_pipeline_t* pipeline, uint32_t max_calls);
TK_NODISCARD tk_error_code_t tk_b200_pipeline_stage_times(tk_b200_pipeline_t* pipeline, double ms_sum[3], uint32_t* out_calls);

void*  tk_b200_pipeline_tensor_device_ptr(tk_b200_pipeline_t* pipeline);
size_t tk_b200_pipeline_frame_bytes(const tk_b200_pipeline_t* pipeline);
size_t tk_b200_pipeline_tensor_bytes(const tk_b200_pipeline_t* pipeline);   /* per frame */

/* ------------------------------------------------------------------------ */
/* Multi-device harness: ONE process, one host thread + one        */
/* pipeline (own streams) per device, frames sharded by contiguous block        */
/* [g*n/G, (g+1)*n/G), no exchange step and no collective: every device writes  */
/* its block's outputs straight into the caller's host arrays.  The reference    */
/* has one device per dispatcher and no multi-device form                        */
/* (src/gpu/cuda/tk_cuda_dispatch.h:47-49, tk_cuda_dispatch.cu:163-247).         */
/* ------------------------------------------------------------------------ */
typedef struct tk_b200_shard_s tk_b200_shard_t;

/* device_ids[num_devices] (an id may repeat); config as tk_b200_pipeline_create with device_id ignored and max_batch =
 * the WHOLE job's batch.  The host threads are bound to the CPUs of their GPU's NUMA node when sysfs names them. */
TK_NODISCARD tk_error_code_t tk_b200_shard_create(tk_b200_shard_t** out_shard, const int* device_ids, uint32_t num_devices,
                                                  const tk_b200_pipeline_config_t* config);
void tk_b200_shard_destroy(tk_b200_shard_t** shard);
uint32_t tk_b200_shard_num_devices(const tk_b200_shard_t* shard);
void tk_b200_shard_block(uint32_t num_frames, uint32_t device_index, uint32_t num_devices, uint32_t* out_first, uint32_t* out_count);
TK_NODISCARD tk_error_code_t tk_b200_shard_update_thresholds(tk_b200_shard_t* shard, float confidence_threshold, float iou_threshold);
/* Buffers of the whole job, laid out as for tk_b200_pipeline_process_host; blocks until every device is done. */
TK_NODISCARD tk_error_code_t tk_b200_shard_process_host(tk_b200_shard_t* shard, const void* h_frames, const float* h_logits,
                                                        uint32_t num_frames, void* h_tensor, tk_b200_candidate_t* h_detections,
                                                        uint32_t* h_num_detections);
TK_NODISCARD tk_error_code_t tk_b200_shard_candidate_overflow(tk_b200_shard_t* shard, uint32_t* out_frames, uint32_t* out_max_count);
/* Final host gather: padded per-frame lists -> flat list in frame order + offsets[n + 1] (prefix sum of the counts).
 * out_flat may be NULL to size it. */
TK_NODISCARD tk_error_code_t tk_b200_shard_gather(const tk_b200_candidate_t* h_detections, const uint32_t* h_num_detections,
                                                  uint32_t num_frames, uint32_t max_detections, tk_b200_candidate_t* out_flat,
                                                  uint64_t* out_offsets);

/* Single frame, reference result type: completes postprocess_detections
 * (src/vision/tk_object_detector.c:303-368) for a HOST logit tensor
 * [num_proposals][5+num_classes].  *out_results is malloc'd (free with
 * tk_object_detector_free_results, src/vision/tk_object_detector.h:157); label
 * points into class_labels (tk_object_detector.h:50-52) or is NULL. */
TK_NODISCARD tk_error_code_t tk_b200_postprocess_detections(
    const float* h_logits, uint32_t num_proposals, uint32_t num_classes, tk_b200_activation_e activation,
    float confidence_threshold, float iou_threshold, uint32_t max_detections,
    const char* const* class_labels, size_t class_count,
    uint32_t frame_width, uint32_t frame_height, uint32_t input_width, uint32_t input_height,
    tk_detection_result_t** out_results, size_t* out_result_count);

/* The detector network behind tk_object_detector_detect, as a device-side hook
 * (the reference copies the tensor to the host around
 * OrtRun, src/vision/tk_object_detector.c:266-269,286-294).  Called on the
 * detector's device with
 *   d_input  f32 [num_frames][3][input_height][input_width]  (ready on `stream`)
 *   d_output f32 [num_frames][num_proposals][5 + class_count] (to be written on `stream`)
 * and must only ENQUEUE work on `stream`.  A non-zero return aborts detect()
 * with TK_ERROR_INFERENCE_FAILED. */
typedef tk_error_code_t (*tk_b200_inference_fn)(void* user, const float* d_input, uint32_t num_frames,
                                                uint32_t input_width, uint32_t input_height, float* d_output,
                                                uint32_t num_proposals, uint32_t row_length, cudaStream_t stream);
TK_NODISCARD tk_error_code_t tk_b200_object_detector_bind_inference(
    tk_object_detector_t* detector, tk_b200_inference_fn fn, void* user, uint32_t num_proposals,
    tk_b200_activation_e activation);
/* Batched detect: num_frames frames of ONE geometry and format; out_results[i]
 * is malloc'd per frame (free each with tk_object_detector_free_results). */
TK_NODISCARD tk_error_code_t tk_b200_object_detector_detect_batch(
    tk_object_detector_t* detector, const tk_video_frame_t* frames, uint32_t num_frames,
    tk_detection_result_t** out_results, size_t* out_result_counts);

/* Handle accessors, so tk_kernels_* launchers can run on a dispatcher's buffers
 * and compute stream. */
void*        tk_b200_buffer_device_ptr(tk_gpu_buffer_t buffer);
size_t       tk_b200_buffer_size(tk_gpu_buffer_t buffer);
cudaStream_t tk_b200_dispatch_stream(tk_cuda_dispatcher_t* dispatcher);
int          tk_b200_dispatch_device(tk_cuda_dispatcher_t* dispatcher);

/* The host-pointer entry points (tk_preprocessor_resize_and_normalize_to_chw, tk_classify_*,
 * tk_b200_postprocess_detections) keep one stream and grow-only device buffers PER CALLING THREAD (the reference's CPU
 * functions are re-entrant; thread-local scratch keeps that without a lock).  They are freed when the thread exits;
 * this frees them now. */
void tk_b200_thread_release(void);

/* Pinned (page-locked, portable) host memory for the host-buffer entry points. */
TK_NODISCARD tk_error_code_t tk_b200_host_alloc(void** out_ptr, size_t size_bytes);
void tk_b200_host_free(void* ptr);

/* Library self-description (no GPU needed). */
const char* tk_b200_version(void);
unsigned long long tk_b200_launch_count(void);
/* L2 fetch-granularity hint (cudaLimitMaxL2FetchGranularity) of the current device's
 * context: bytes = 32 / 64 / 128 sets it, 0 only queries; returns the value in effect. */
int tk_b200_l2_fetch_granularity(int bytes);   /* kernels enqueued by this library so far */
int tk_b200_device_count(void);
/* Debug: register a zeroed DEVICE buffer; the pre-process and post-process kernels then append one 64-byte record per
 * CTA / per frame ({kernel id << 32 | block, smid << 32 | frame, six %globaltimer stamps}; word 0 of the buffer counts
 * the records, record i starts at byte 64 (i + 1)).  NULL or 0 bytes turns it off.  Process-wide, not for production. */
void tk_b200_debug_trace(void* d_buf, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* TK_B200_VISION_H */
