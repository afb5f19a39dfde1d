// color.cu — per-detection attribute kernels: dominant colour and door state.
//
// Replaces  tk_classify_dominant_color (src/vision/tk_attribute_classifier.c:40-97,
//             rgb_to_hsv :19-36) and tk_classify_door_state (:100-157), which the
//             pipeline runs serially per detection on the CPU
//             (src/vision/tk_vision_pipeline.c:451-488).
//
// One CTA per detection box; threads stride over the clipped box, histogram in
// registers, warp-shuffle + shared-memory reduction, first-max argmax (strict `>`,
// lowest bin wins ties, :83-88).  The reference's mixed float/double hue
// arithmetic is reproduced operation by operation so that the bin of every pixel
// — and therefore the argmax — is bit-identical.
#include "tk_common.cuh"

namespace tkb {

static constexpr int kColThreads = 256;

// src/vision/tk_attribute_classifier.c:19-36 and :66-79.  Returns 0..8, or -1
// when no bin is incremented.
__device__ __forceinline__ int color_bin(uint32_t R, uint32_t G, uint32_t B) {
    const float r = __fdiv_rn((float)R, 255.0f), g = __fdiv_rn((float)G, 255.0f), b = __fdiv_rn((float)B, 255.0f);
    const float mx = fmaxf(r, fmaxf(g, b)), mn = fminf(r, fminf(g, b));
    const float delta = __fsub_rn(mx, mn);
    float s = 0.0f, h = 0.0f;
    if (mx > 0.0f) {
        s = __fdiv_rn(delta, mx);
        if (r >= mx) h = __fdiv_rn(__fsub_rn(g, b), delta);
        else if (g >= mx) h = (float)__dadd_rn(2.0, (double)__fdiv_rn(__fsub_rn(b, r), delta));
        else h = (float)__dadd_rn(4.0, (double)__fdiv_rn(__fsub_rn(r, g), delta));
        h = (float)__dmul_rn((double)h, 60.0);
        if ((double)h < 0.0) h = (float)__dadd_rn((double)h, 360.0);
    }
    if ((double)s < 0.1) {
        if ((double)mx < 0.1) return 6;
        if ((double)mx > 0.9) return 7;
        return 8;
    }
    if (h < 30.0f || h >= 330.0f) return 0;
    if (h < 90.0f) return 1;
    if (h < 150.0f) return 2;
    if (h < 210.0f) return 3;
    if (h < 270.0f) return 4;
    if (h < 330.0f) return 5;
    return -1;
}

__global__ void __launch_bounds__(kColThreads) dominant_color_kernel(
    const uint8_t* __restrict__ frame, uint32_t W, uint32_t H, uint32_t stride, const tk_rect_t* __restrict__ boxes,
    uint32_t* __restrict__ ids, unsigned long long* __restrict__ bins_out) {
    __shared__ unsigned int sbin[9];
    __shared__ unsigned long long tot[9];
    const tk_rect_t bb = boxes[blockIdx.x];
    // clip (the reference skips out-of-frame pixels one by one, :56); 64-bit ends: x+w may exceed int
    const long long x0 = max((long long)bb.x, 0ll), y0 = max((long long)bb.y, 0ll);
    const long long x1 = min((long long)bb.x + bb.w, (long long)W), y1 = min((long long)bb.y + bb.h, (long long)H);
    if (threadIdx.x < 9) { sbin[threadIdx.x] = 0; tot[threadIdx.x] = 0; }
    __syncthreads();
    const long long bw = x1 - x0, bh = y1 - y0;
    if (bw > 0 && bh > 0) {
        const unsigned long long n = (unsigned long long)bw * (unsigned long long)bh;
        // chunks of <= 2^24 pixels keep the 32-bit shared counters exact
        for (unsigned long long c0 = 0; c0 < n; c0 += (1ull << 24)) {
            const unsigned long long c1 = min(n, c0 + (1ull << 24));
            uint32_t cnt[9];
#pragma unroll
            for (int i = 0; i < 9; ++i) cnt[i] = 0;
            for (unsigned long long i = c0 + threadIdx.x; i < c1; i += kColThreads) {
                const uint32_t yy = (uint32_t)(i / (unsigned long long)bw), xx = (uint32_t)(i % (unsigned long long)bw);
                const uint8_t* p = frame + (size_t)(y0 + yy) * stride + (size_t)(x0 + xx) * 3;
                const int kbin = color_bin(p[0], p[1], p[2]);
#pragma unroll
                for (int q = 0; q < 9; ++q) cnt[q] += (kbin == q);
            }
#pragma unroll
            for (int q = 0; q < 9; ++q) {
                uint32_t v = cnt[q];
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
                if ((threadIdx.x & 31) == 0 && v) atomicAdd(&sbin[q], v);
            }
            __syncthreads();
            if (threadIdx.x < 9) { tot[threadIdx.x] += sbin[threadIdx.x]; sbin[threadIdx.x] = 0; }
            __syncthreads();
        }
    }
    if (threadIdx.x == 0) {
        int best = 0;
        for (int i = 1; i < 9; ++i) if (tot[i] > tot[best]) best = i;
        ids[blockIdx.x] = (uint32_t)best;
        if (bins_out) for (int i = 0; i < 9; ++i) bins_out[(size_t)blockIdx.x * 9 + i] = tot[i];
    }
}

// src/vision/tk_attribute_classifier.c:119-147.  The reference does not clip the
// box (it reads outside the frame); here pixels whose 3-row column leaves the
// frame are skipped.  edges[] gets the strong-edge count, states[] 1 = "closed".
__device__ __forceinline__ int gray_u8(const uint8_t* p) { return (int)(((uint32_t)p[0] * 299u + (uint32_t)p[1] * 587u + (uint32_t)p[2] * 114u) / 1000u); }

__global__ void __launch_bounds__(kColThreads) door_state_kernel(
    const uint8_t* __restrict__ frame, uint32_t W, uint32_t H, uint32_t stride, const tk_rect_t* __restrict__ boxes,
    uint32_t* __restrict__ states, int32_t* __restrict__ edges) {
    __shared__ unsigned int total;
    const tk_rect_t bb = boxes[blockIdx.x];
    if (threadIdx.x == 0) total = 0;
    __syncthreads();
    const long long x0 = max((long long)bb.x + 1, 0ll), y0 = max((long long)bb.y + 1, 1ll);
    const long long x1 = min((long long)bb.x + bb.w - 1, (long long)W), y1 = min((long long)bb.y + bb.h - 1, (long long)H - 1);
    const long long bw = x1 - x0, bh = y1 - y0;
    uint32_t cnt = 0;
    if (bw > 0 && bh > 0) {
        const unsigned long long n = (unsigned long long)bw * (unsigned long long)bh;
        for (unsigned long long i = threadIdx.x; i < n; i += kColThreads) {
            const uint32_t yy = (uint32_t)(i / (unsigned long long)bw), xx = (uint32_t)(i % (unsigned long long)bw);
            const uint8_t* above = frame + (size_t)(y0 + yy - 1) * stride + (size_t)(x0 + xx) * 3;
            const uint8_t* below = above + 2 * (size_t)stride;
            cnt += abs(gray_u8(above) - gray_u8(below)) > 100;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
    if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&total, cnt);
    __syncthreads();
    if (threadIdx.x == 0) {
        const float density = __fdiv_rn((float)(int)total, (float)(int)((unsigned)bb.w * (unsigned)bb.h));   // :141
        states[blockIdx.x] = ((double)density > 0.1) ? 1u : 0u;
        if (edges) edges[blockIdx.x] = (int32_t)total;
    }
}

}  // namespace tkb

using namespace tkb;

static const char* const kColorNames[9] = {"red", "yellow", "green", "cyan", "blue", "magenta", "black", "white", "gray"};

extern "C" const char* tk_b200_color_name(uint32_t id) { return id < 9 ? kColorNames[id] : "unknown"; }

extern "C" tk_error_code_t tk_kernels_dominant_color(const tk_b200_color_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_frame || !p->d_color_ids || (!p->d_boxes && p->num_boxes)) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_boxes == 0) return TK_SUCCESS;
    if (p->width == 0 || p->height == 0) return TK_ERROR_INVALID_DIMENSIONS;
    if (!device_info()) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    const uint32_t stride = p->stride_bytes ? p->stride_bytes : p->width * 3;
    dominant_color_kernel<<<p->num_boxes, kColThreads, 0, stream>>>(p->d_frame, p->width, p->height, stride, p->d_boxes,
                                                                   p->d_color_ids, p->d_bins);
    return launch_status("dominant_color_kernel");
}

extern "C" tk_error_code_t tk_kernels_door_state(const tk_b200_door_params_t* p, cudaStream_t stream) {
    if (!p || !p->d_frame || !p->d_states || (!p->d_boxes && p->num_boxes)) return TK_ERROR_INVALID_ARGUMENT;
    if (p->num_boxes == 0) return TK_SUCCESS;
    if (p->width == 0 || p->height == 0) return TK_ERROR_INVALID_DIMENSIONS;
    if (!device_info()) return TK_ERROR_GPU_DEVICE_NOT_FOUND;
    const uint32_t stride = p->stride_bytes ? p->stride_bytes : p->width * 3;
    door_state_kernel<<<p->num_boxes, kColThreads, 0, stream>>>(p->d_frame, p->width, p->height, stride, p->d_boxes,
                                                               p->d_states, p->d_edge_pixels);
    return launch_status("door_state_kernel");
}
