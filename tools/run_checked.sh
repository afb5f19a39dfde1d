#!/bin/bash
# The CHECKED build (device-side assertions + ring-content verification, csrc/tk_common.cuh) over the every-variant
# cases, the ring stress tests and the full-size geometries — what stands in for compute-sanitizer, which is closed on
# this GPU pool.  Run on the GPU box:  bash tools/run_checked.sh   -> gpurun_out/r2_checked.log
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
[ -f trackiellm_b200/libtk_b200_checked.so ] || make -s -C trackiellm_b200/csrc checked
export TK_B200_LIB=$PWD/trackiellm_b200/libtk_b200_checked.so
python -m pytest -q -m gpu -p no:cacheprovider \
    tests/test_gpu_sanitizer_cases.py tests/test_gpu_post_fused.py tests/test_gpu_postprocess.py tests/test_gpu_pipeline.py \
    "tests/test_gpu_preprocess.py::test_ring_stress_many_rows_per_cta" "tests/test_gpu_preprocess.py::test_random_geometries_all_formats_bitexact" \
    "tests/test_gpu_preprocess.py::test_fp16_vs_oracle" "tests/test_gpu_preprocess.py::test_i420" "tests/test_gpu_preprocess.py::test_letterbox" \
    "tests/test_gpu_preprocess.py::test_stride_and_unaligned_fallback" "tests/test_gpu_preprocess.py::test_dynamic_row_share_extremes_bitexact" "tests/test_gpu_preprocess.py::test_concurrent_launches_on_two_streams_take_different_chunk_counters" "tests/test_gpu_baseline_configs.py::test_cfg1_64_frames_1080p_fp16" \
    > gpurun_out/r2_checked.log 2>&1
echo "checked build: exit $? ; $(tail -1 gpurun_out/r2_checked.log)"
grep -c "TK_B200_CHECKED" gpurun_out/r2_checked.log
