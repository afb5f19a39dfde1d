#!/bin/bash
# compute-sanitizer over the small every-variant GPU cases (tests/test_gpu_sanitizer_cases.py), one tool per pass.
# Run on the GPU box:  bash tools/run_sanitizer.sh [tool ...]   (default: memcheck racecheck synccheck initcheck)
# Logs: gpurun_out/san_<tool>.log (sanitizer) and gpurun_out/san_<tool>.pytest.log; summary lines on stdout.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TOOLS=${@:-memcheck racecheck synccheck initcheck}
SEL=${SAN_SELECT:-tests/test_gpu_sanitizer_cases.py}
for tool in $TOOLS; do
  extra=""
  [ "$tool" = racecheck ] && extra="--racecheck-report all"
  [ "$tool" = initcheck ] && extra="--track-unused-memory no"
  t0=$(date +%s)
  timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool $extra --print-limit 40 --error-exitcode 86 \
      --log-file gpurun_out/san_$tool.log \
      python -m pytest $SEL -q -m gpu -p no:cacheprovider > gpurun_out/san_$tool.pytest.log 2>&1
  rc=$?
  t1=$(date +%s)
  echo "== $tool: exit $rc, $((t1 - t0)) s; $(tail -1 gpurun_out/san_$tool.pytest.log)"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard" gpurun_out/san_$tool.log | sort | uniq -c | sort -rn | head -12
done
