O=gpurun_out/d5; mkdir -p $O
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -3 $O/smoke.log | cut -c1-600
timeout 1500 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" >> $O/gpu_tests.log; tail -4 $O/gpu_tests.log
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; tail -c 500 $O/bench_ref.json
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_cfg3.json 2> $O/bench_cfg3.err; tail -c 300 $O/bench_cfg3.json
