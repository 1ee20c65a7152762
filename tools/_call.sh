O=gpurun_out/d11; mkdir -p $O
timeout 1500 python -m pytest tests -x -q -m gpu > $O/gpu_tests.log 2>&1; echo "gpu tests rc=$?" >> $O/gpu_tests.log; tail -4 $O/gpu_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -1 $O/smoke.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; python -c "
import json; d=json.loads(open('$O/bench_default.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value']/1e6, d['e2e']['ms_per_step'], d['gpu_launches_per_step'], d['roofline']['kernel'][:30], d['roofline']['frac'], d['path']['frac_of_sustained_peak'], d['clocks'], d['cpu_baseline']['value'])"
