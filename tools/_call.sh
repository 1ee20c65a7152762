timeout 900 python tools/parity_report.py > gpurun_out/parity_report.log 2>&1; tail -25 gpurun_out/parity_report.log
