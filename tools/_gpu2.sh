set -x
timeout 600 python tools/exp_coresident.py r2_cores_exact exact > gpurun_out/r2_cores_exact.txt 2>&1; cat gpurun_out/r2_cores_exact.txt
timeout 600 python tools/exp_coresident.py r2_cores_half half > gpurun_out/r2_cores_half.txt 2>&1; cat gpurun_out/r2_cores_half.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2_gpu_tests_a.log 2>&1; tail -5 gpurun_out/r2_gpu_tests_a.log
timeout 900 python bench.py --steps 50 --warmup 5 --fuse 1 > gpurun_out/r2_bench_new.json 2> gpurun_out/r2_bench_new.err; tail -c 1500 gpurun_out/r2_bench_new.json; tail -5 gpurun_out/r2_bench_new.err
timeout 600 ncu --set full --import-source on --clock-control none -k regex:preprocess_kernel -s 10 -c 1 -f -o gpurun_out/r2_pre_exact python tools/exp_pre.py > gpurun_out/r2_ncu_pre.log 2>&1; tail -3 gpurun_out/r2_ncu_pre.log
timeout 600 ncu --set full --import-source on --clock-control none -k regex:post_kernel -s 2 -c 1 -f -o gpurun_out/r2_post_alone python tools/exp_coresident.py r2_ncu_dummy exact > gpurun_out/r2_ncu_post.log 2>&1; tail -3 gpurun_out/r2_ncu_post.log
ls -la gpurun_out/*.ncu-rep
