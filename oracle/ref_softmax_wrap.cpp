// C entry point around the reference's own softmax_cpu
// (tests/tk_gpu_softmax_test.cpp:24-51), which is compiled from the reference
// tree in place by oracle/Makefile with its main() renamed.  TEST INFRASTRUCTURE.
#include <vector>
void softmax_cpu(const std::vector<float>& input, std::vector<float>& output, int rows, int cols);
extern "C" void ref_softmax_cpu(const float* in, float* out, int rows, int cols) {
    std::vector<float> vi(in, in + (size_t)rows * cols), vo;
    softmax_cpu(vi, vo, rows, cols);
    for (size_t i = 0; i < vo.size(); ++i) out[i] = vo[i];
}
