/*
 * tk_oracle.h — CPU oracle for the TrackieLLM vision hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under trackiellm_b200/ may include, link
 * or call this; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs do, and only as the checker or the CPU
 * baseline.
 *
 * Parity status of each function:
 *   PINNED    oracle_preprocess_rgb8, oracle_dominant_color, oracle_door_state
 *             — checked bit-for-bit against the reference's own C files
 *             compiled into oracle/_ref/libtkref.so (oracle/Makefile) and
 *             against the reference's known-answer tests
 *             (tests/tk_attribute_classifier_test.c:21-93);
 *             oracle_softmax_rows — checked against the reference's
 *             softmax_cpu (tests/tk_gpu_softmax_test.cpp:24-51, 4x256 KAT).
 *   UNPINNED  oracle_decode, oracle_nms_classwise, oracle_restore_coords,
 *             oracle_i420_to_rgb8, letterbox mode — the reference holds only
 *             stubs/fragments for these (src/vision/tk_object_detector.c:303-375
 *             cannot compile and elides class selection and NMS); the
 *             restatement follows every fragment the reference does pin and
 *             says "parity unpinned" for the rest.
 */
#ifndef TK_ORACLE_H
#define TK_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int32_t  index;        /* proposal row within the frame */
    int32_t  x, y, w, h;   /* tk_rect_t, truncated ints      */
    float    confidence;
    uint32_t class_id;
} oracle_candidate_t;

enum { ORACLE_ACT_RAW = 0, ORACLE_ACT_SIGMOID = 1, ORACLE_ACT_SOFTMAX = 2 };
enum { ORACLE_NORM_AVX = 0, ORACLE_NORM_SCALAR = 1 };
enum { ORACLE_RESIZE_STRETCH = 0, ORACLE_RESIZE_LETTERBOX = 1 };

/* (a2) tk_preprocessor_resize_and_normalize_to_chw restated; honours stride. */
int oracle_preprocess_rgb8(const uint8_t* src, uint32_t src_w, uint32_t src_h,
                           uint32_t src_stride, float* out, uint32_t dst_w,
                           uint32_t dst_h, const float mean[3], const float std_dev[3],
                           int norm_variant, int resize_mode, float pad_value);

/* I420 -> packed RGB8 (BT.601 full range, 16.16 fixed point, nearest chroma). */
void oracle_i420_to_rgb8(const uint8_t* i420, uint32_t w, uint32_t h, uint8_t* rgb);

/* fp32 -> IEEE binary16 bits, round to nearest even. */
uint16_t oracle_f32_to_f16(float f);
void oracle_f32_to_f16_array(const float* in, uint16_t* out, size_t n);

/* (a5) decode; returns the TOTAL number of rows passing the threshold,
 * writes the first min(total, cap) candidates in ascending row order. */
uint32_t oracle_decode(const float* logits, uint32_t num_proposals, uint32_t num_classes,
                       float conf_threshold, int activation,
                       oracle_candidate_t* out, uint32_t cap);

/* (a6) IoU on integer rects. */
float oracle_iou(const oracle_candidate_t* a, const oracle_candidate_t* b);

/* (a7) class-wise greedy NMS. Returns number kept (<= max_detections). */
uint32_t oracle_nms_classwise(const oracle_candidate_t* cands, uint32_t count,
                              float iou_threshold, uint32_t max_detections,
                              oracle_candidate_t* kept);

/* (a9) model-input px -> source-frame px (stretch inverse). */
void oracle_restore_coords(oracle_candidate_t* dets, uint32_t n, uint32_t src_w,
                           uint32_t src_h, uint32_t in_w, uint32_t in_h);

/* (a10) row softmax, reference arithmetic order; argmax = first max. */
void oracle_softmax_rows(const float* in, float* out, uint32_t rows, uint32_t cols,
                         int use_det_exp);
void oracle_argmax_rows(const float* in, uint32_t* out, uint32_t rows, uint32_t cols);

/* (a11) dominant colour: writes the 9 bin counts and returns the argmax bin. */
int oracle_dominant_color(const uint8_t* rgb, uint32_t w, uint32_t h, uint32_t stride,
                          int bx, int by, int bw, int bh, int64_t bins[9]);
/* door-state edge density; returns 1 = "closed", 0 = "open". */
int oracle_door_state(const uint8_t* rgb, uint32_t w, uint32_t h, uint32_t stride,
                      int bx, int by, int bw, int bh, int32_t* edge_pixels);

/* (f1) MiDaS inverse depth -> metric, src/vision/tk_depth_midas.c:473-499 (unpinned by execution). */
void oracle_inverse_depth_to_metric(const float* in, float* out, size_t n, float min_depth, float max_depth);
/* (f2) per-detection distance, src/vision/src/object_analysis.rs:227-279; scratch: depth_w*depth_h floats. */
float oracle_detection_distance(int bx, int by, int bw, int bh, const float* depth, uint32_t depth_w,
                                uint32_t depth_h, uint32_t frame_w, uint32_t frame_h, float* scratch);

float oracle_det_expf(float x);
float oracle_det_sigmoidf(float x);

#ifdef __cplusplus
}
#endif
#endif
