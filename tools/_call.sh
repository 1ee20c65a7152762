#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/t1_n8.json 2> gpurun_out/t1_n8.err; echo "n8 rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/t1_n1.json 2> gpurun_out/t1_n1.err; echo "n1 rc=$?"
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --workload aishell_train_step_B32_T500_U50 > gpurun_out/t1_n8_train.json 2> gpurun_out/t1_n8_train.err; echo "n8 train rc=$?"
for f in t1_n8 t1_n1 t1_n8_train; do python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
    print("$f", round(l["ms_per_step"],3), round(l["value"]/1e6,2), "e2e", round(l["e2e"]["ms_per_step"],3), l["path"].get("step_ms_per_rank"), l["path"].get("allreduce_ms_per_rank"), l["clocks"]["sm_mhz"])
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/$f.err").read()[-1500:])
PY
done
