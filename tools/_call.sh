#!/bin/bash
mkdir -p gpurun_out
KERNEL=joint_fwd_store MODE=2 DBGS=0,4 timeout 900 python tools/bwd_debug_matrix.py - RNNT_B200_EBLOCK=1 - RNNT_B200_EBLOCK=1 > gpurun_out/g5_matrix.txt 2>&1
cat gpurun_out/g5_matrix.txt
