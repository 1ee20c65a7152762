#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/v1_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/v1_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/v1_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/v1_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/v1_bench.json 2> gpurun_out/v1_bench.err; echo "bench rc=$?"; python - <<PY
import json
l=json.loads(open("gpurun_out/v1_bench.json").read().strip().splitlines()[-1])
print(round(l["ms_per_step"],3), round(l["value"]/1e6,2), "e2e", round(l["e2e"]["ms_per_step"],3), l["roofline"]["frac"], l["path"]["frac_of_sustained_peak"], l["path"]["kernel_ms_one_step"], l["clocks"])
PY
