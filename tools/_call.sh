set -x
mkdir -p gpurun_out/c1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1/smi.txt
timeout 300 python -m pytest tests/test_fused_gpu.py -q -m gpu -k "dh_gemm or dpre_reduce or dw2_gemm" -x > gpurun_out/c1/kern_tests.log 2>&1
echo "kern tests rc=$?" >> gpurun_out/c1/kern_tests.log
tail -5 gpurun_out/c1/kern_tests.log
timeout 200 python tools/dh_gemm_bench.py > gpurun_out/c1/dh_bench.log 2>&1; tail -3 gpurun_out/c1/dh_bench.log
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/c1/gpu_tests.log 2>&1
echo "gpu tests rc=$?" >> gpurun_out/c1/gpu_tests.log
tail -15 gpurun_out/c1/gpu_tests.log
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/c1/bench_cfg3.json 2> gpurun_out/c1/bench_cfg3.err; tail -c 3000 gpurun_out/c1/bench_cfg3.json
timeout 300 python bench.py --steps 10 --warmup 3 --workload aishell_long_B16_T1500_U200 --no-cpu-baseline > gpurun_out/c1/bench_long.json 2> gpurun_out/c1/bench_long.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload aishell_train_step_B32_T500_U50 --no-cpu-baseline > gpurun_out/c1/bench_train.json 2> gpurun_out/c1/bench_train.err; tail -c 1500 gpurun_out/c1/bench_train.json
