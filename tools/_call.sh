#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py -x -q -m gpu -k "spread or synthetic or golden or stored or config or graphed or headline" > gpurun_out/x1_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/x1_tests.log
