O=gpurun_out/d2; mkdir -p $O
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > $O/zz_alone.log 2>&1; echo "alone rc=$?" >> $O/zz_alone.log; tail -4 $O/zz_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" >> $O/gpu_tests.log; tail -4 $O/gpu_tests.log
