#!/bin/bash
mkdir -p gpurun_out
timeout 900 python tools/parity_report.py > gpurun_out/w1_parity.txt 2>&1; echo "rc=$?"; tail -25 gpurun_out/w1_parity.txt | cut -c1-700
