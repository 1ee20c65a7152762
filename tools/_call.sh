#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
for v in nobal1 bal1 nobal2 bal2; do
  fl=""; case $v in bal*) fl="--balance";; esac
  timeout 240 $TR bench.py --gpus 8 --steps 20 --warmup 5 $fl > gpurun_out/u1_n8_$v.json 2> gpurun_out/u1_n8_$v.err; echo "$v rc=$?"
done
for f in nobal1 bal1 nobal2 bal2; do python - <<PY
import json
try:
    l=json.loads(open("gpurun_out/u1_n8_$f.json").read().strip().splitlines()[-1])
    print("$f", round(l["ms_per_step"],3), round(l["value"]/1e6,2), round(l["e2e"]["ms_per_step"],3), l["path"].get("speed_weighted_sharding"), l["path"].get("allreduce_ms_per_rank"))
except Exception as e:
    print("$f FAILED", e); print(open("gpurun_out/u1_n8_$f.err").read()[-1500:])
PY
done
