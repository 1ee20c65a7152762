#!/bin/bash
mkdir -p gpurun_out
RNNT_B200_PAIR_SCHED=2phase timeout 900 python -m pytest tests/test_fused_gpu.py -x -q -m gpu -k "synthetic or golden or stored or config or chunks or headline" > gpurun_out/r1_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r1_tests.log
run() { # name sched frac
  RNNT_B200_PAIR_SCHED=$2 RNNT_B200_PAIR_F=$3 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r1_$1.json 2> gpurun_out/r1_$1.err
  python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/r1_$1.json").read().strip().splitlines()[-1])
    k=l["path"]["kernel_ms_one_step"]
    print("$1 $2 $3", round(l["ms_per_step"],3), "e2e", round(l["e2e"]["ms_per_step"],3), "fwd", k["joint_fwd_store"], "pair", k["dh_dw2_pair"], l["clocks"]["sm_mhz"])
except Exception as e:
    print("$1 FAILED", e); print(open("gpurun_out/r1_$1.err").read()[-1500:])
PY
}
run s1 side 21/50
run a1 2phase 21/50
run b1 2phase 19/50
run c1 2phase 23/50
run s2 side 21/50
run a2 2phase 21/50
run b2 2phase 19/50
run c2 2phase 23/50
