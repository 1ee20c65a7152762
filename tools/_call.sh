#!/bin/bash
# scratch GPU script of the current call
mkdir -p gpurun_out
timeout 300 python tools/profile_range.py step > gpurun_out/k1_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --replay-mode app-range --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/k1_range_step.csv python tools/profile_range.py step > gpurun_out/k1_ncu.log 2>&1; echo "ncu rc=$?"
cat gpurun_out/k1_range_step.csv | tail -8
