#!/bin/bash
# scratch GPU script of the current call
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/fin_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/fin_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/fin_smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py > gpurun_out/fin_bench.json 2> gpurun_out/fin_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/fin_ref.json 2> gpurun_out/fin_ref.err; echo "ref rc=$?"
python - <<PY
import json
for f in ["fin_bench","fin_ref"]:
    l=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, l.get("impl"), round(l["ms_per_step"],3), round(l["value"],1), "e2e", round(l["e2e"]["value"],1), l.get("clocks"))
PY
