#!/bin/bash
# scratch GPU script of the current call
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fused_gpu.py -x -q -m gpu -k "synthetic or golden or row_factors or stored or either_sign or config or deterministic or chunks or headline or graphed or hypothesis" > gpurun_out/f1_tests.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/f1_tests.log
for v in stored recompute stored2 recompute2; do
  RNNT_B200_BWD=${v%2} timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/f1_$v.json 2> gpurun_out/f1_$v.err; echo "$v rc=$?"
done
for f in stored recompute stored2 recompute2; do python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/f1_$f.json").read().strip().splitlines()[-1])
    print("$f", round(l["ms_per_step"],3), round(l["value"]/1e6,2), "e2e", round(l["e2e"]["ms_per_step"],3), "loss", l["loss"], l["path"]["kernel_ms_one_step"], l["clocks"]["sm_mhz"])
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/f1_$f.err").read()[-1500:])
PY
done
