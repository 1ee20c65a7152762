#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python tools/soak_fused.py 300 7 stored > gpurun_out/z2_soak_stored.txt 2>&1; echo "soak stored rc=$?"; grep -E "^worst|^soak" gpurun_out/z2_soak_stored.txt
RNNT_B200_BWD=recompute timeout 1200 python tools/soak_fused.py 300 7 stored > gpurun_out/z2_soak_recompute.txt 2>&1; echo "soak recompute rc=$?"; grep -E "^worst|^soak" gpurun_out/z2_soak_recompute.txt
