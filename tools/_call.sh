O=gpurun_out/d14; mkdir -p $O
timeout 900 python -m pytest tests/test_fused_gpu.py -q -m gpu -k "hypothesis" > $O/hyp.log 2>&1; tail -30 $O/hyp.log | cut -c1-400
