#!/bin/bash
mkdir -p gpurun_out
echo skip tests
for v in p2a off1 p3a p2b off2 p3b; do
  case $v in p2*) po=2;; p3*) po=3;; off*) po=0;; esac
  RNNT_B200_PATCH_OVERLAP=$po timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/m1_$v.json 2> gpurun_out/m1_$v.err; echo "$v rc=$?"
done
for f in p2a off1 p3a p2b off2 p3b; do python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/m1_$f.json").read().strip().splitlines()[-1])
    print("$f", round(l["ms_per_step"],3), round(l["value"]/1e6,2), "e2e", round(l["e2e"]["ms_per_step"],3), l["path"]["kernel_ms_one_step"], l["clocks"]["sm_mhz"])
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/m1_$f.err").read()[-1500:])
PY
done
