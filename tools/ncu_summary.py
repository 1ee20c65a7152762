"""Key metrics per kernel from `ncu -i X.ncu-rep --page raw --csv` (stdin or file arg)."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = [r'^gpu__time_duration.sum$', r'^sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_(active|elapsed)$',
        r'^sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active$', r'^sm__throughput.avg.pct',
        r'^dram__bytes_read.sum$', r'^dram__bytes_write.sum$', r'^dram__throughput.avg.pct_of_peak_sustained_elapsed$',
        r'^gpu__dram_throughput.avg.pct', r'^lts__throughput.avg.pct', r'^l1tex__throughput.avg.pct',
        r'^sm__warps_active.avg.pct', r'^launch__registers_per_thread$', r'^launch__grid_size$', r'^launch__block_size$',
        r'^sm__cycles_elapsed.avg$', r'^sm__cycles_elapsed.avg.per_second$', r'^smsp__issue_active.avg.pct',
        r'^launch__cluster', r'^smsp__pcsamp_warps_issue_stalled_(long_scoreboard|barrier|wait|short_scoreboard|math_pipe_throttle|mio_throttle|lg_throttle|selected|sleeping)$']
idx = [i for i, h in enumerate(hdr) if any(re.search(p, h) for p in want)]
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    print("=====", r[ki][:100])
    for i in idx:
        if r[i] not in ("", "n/a", "no data"):
            print(f"  {hdr[i]:78s} {units[i]:16s} {r[i][:24]}")
