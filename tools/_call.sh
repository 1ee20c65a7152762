#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/y1_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/y1_tests.log
timeout 300 python tools/profile_joint.py 32 500 100 > gpurun_out/y1_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/y1_launches.csv python tools/profile_joint.py 32 500 100 > gpurun_out/y1_ncu.log 2>&1; echo "ncu rc=$?"
python tools/ncu_traffic.py gpurun_out/y1_launches.csv gpurun_out/y1_traffic.json > gpurun_out/y1_launches.txt 2>&1; head -12 gpurun_out/y1_launches.txt; tail -1 gpurun_out/y1_launches.txt
