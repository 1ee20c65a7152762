O=gpurun_out/d1; mkdir -p $O
timeout 300 python -m pytest tests/test_fused_gpu.py -q -m gpu -k "dh_gemm or dpre_reduce or dw2_gemm or synthetic or deterministic" > $O/kern_tests.log 2>&1; tail -3 $O/kern_tests.log
timeout 300 python tools/trace_dh.py > $O/trace_dh.log 2>&1; cat $O/trace_dh.log
timeout 200 python tools/dh_gemm_bench.py 3256 > $O/dh_bench.log 2>&1; tail -1 $O/dh_bench.log
timeout 1500 python -m pytest tests -q -m gpu > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" >> $O/gpu_tests.log; tail -4 $O/gpu_tests.log
timeout 400 python bench.py --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; tail -c 400 $O/bench_cfg3.json; tail -3 $O/bench_cfg3.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload aishell_long_B16_T1500_U200 --no-cpu-baseline > $O/bench_long.json 2> $O/bench_long.err
timeout 300 python bench.py --steps 10 --warmup 3 --workload aishell_train_step_B32_T500_U50 --no-cpu-baseline > $O/bench_train.json 2> $O/bench_train.err; tail -c 300 $O/bench_train.json
