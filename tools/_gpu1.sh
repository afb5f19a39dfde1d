set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests/test_gpu_post_fused.py -x -q -m gpu > gpurun_out/r2_fused_tests.log 2>&1; tail -15 gpurun_out/r2_fused_tests.log
timeout 900 python -m pytest tests/test_gpu_sanitizer_cases.py -x -q -m gpu > gpurun_out/r2_san_plain.log 2>&1; tail -5 gpurun_out/r2_san_plain.log
for f in 1 2; do for a in exact half; do
TK_B200_POST_FUSED=$f timeout 600 python bench.py --steps 50 --warmup 5 --arith $a --no-cpu-baseline --e2e-steps 3 > gpurun_out/r2_b_${a}_f$f.json 2> gpurun_out/r2_b_${a}_f$f.err
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2_b_${a}_f$f.json").read().strip().splitlines()[-1])
    print("BENCH $a fused=$f", d["value"], d["ms_per_step"], [ (s["kernel"], round(s["ms_per_launch"]*1e3,1)) for s in d["roofline_stages"]], d["e2e"]["value"])
except Exception as e: print("bench $a $f failed", e); print(open("gpurun_out/r2_b_${a}_f$f.err").read()[-1500:])
PY
done; done
python tools/exp_pre.py > gpurun_out/r2_exp_pre.txt 2>&1; cat gpurun_out/r2_exp_pre.txt
bash tools/run_sanitizer.sh memcheck racecheck synccheck initcheck
