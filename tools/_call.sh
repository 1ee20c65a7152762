#!/bin/bash
# scratch GPU script of the current call
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/h1_tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/h1_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/h1_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/h1_smoke.log | cut -c1-600
timeout 600 python bench.py > gpurun_out/h1_bench.json 2> gpurun_out/h1_bench.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/h1_bench.json
