O=gpurun_out/d8; mkdir -p $O
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
timeout 300 python tools/profile_range.py > $O/plain.log 2>&1 && \
timeout 900 ncu --replay-mode app-range --metrics $M --clock-control none --csv --log-file $O/range_overlap.csv python tools/profile_range.py > $O/ncu_overlap.log 2>&1
tail -3 $O/ncu_overlap.log; cat $O/range_overlap.csv | tail -8
RNNT_B200_BWD_OVERLAP=0 timeout 900 ncu --replay-mode app-range --metrics $M --clock-control none --csv --log-file $O/range_serial.csv python tools/profile_range.py > $O/ncu_serial.log 2>&1
tail -3 $O/ncu_serial.log; cat $O/range_serial.csv | tail -8
