O=gpurun_out/d10; mkdir -p $O
timeout 600 python -m pytest tests/test_fused_gpu.py -q -m gpu -k "graphed" > $O/graph_tests.log 2>&1; tail -15 $O/graph_tests.log | cut -c1-300
timeout 600 python tools/graph_bench.py > $O/graph_bench.log 2>&1; tail -5 $O/graph_bench.log | cut -c1-400
