#!/bin/bash
mkdir -p gpurun_out
timeout 200 python tools/check_hidden_bits.py rnnt-transducer_b200/librnnt_b200_prev.so; timeout 900 python -m pytest tests/test_fused_gpu.py -x -q -m gpu -k "synthetic or golden or stored or config or hypothesis" > gpurun_out/p1_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/p1_tests.log
for v in new1 prev1 new2 prev2 new3 prev3; do
  case $v in prev*) export RNNT_B200_LIB=$PWD/rnnt-transducer_b200/librnnt_b200_prev.so;; new*) unset RNNT_B200_LIB;; esac
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/p1_$v.json 2> gpurun_out/p1_$v.err; echo "$v rc=$?"
done
for f in new1 prev1 new2 prev2 new3 prev3; do python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/p1_$f.json").read().strip().splitlines()[-1])
    k=l["path"]["kernel_ms_one_step"]
    print("$f", round(l["ms_per_step"],3), "e2e", round(l["e2e"]["ms_per_step"],3), "fwd", k["joint_fwd_store"], "hidden", k["joint_hidden"], "patch", k["patch_scale"], "pair", k["dh_dw2_pair"], l["clocks"]["sm_mhz"])
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/p1_$f.err").read()[-1500:])
PY
done
