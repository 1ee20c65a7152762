"""From an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv` log of
tools/profile_joint.py (two identical steps): per-kernel launches, time and DRAM bytes of the SECOND step, and the
step total. Usage: python tools/ncu_traffic.py launches.csv [out.json]"""
import collections, csv, json, sys

def main(path, out=None):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    per_id = collections.OrderedDict()
    for row in csv.DictReader(lines):
        e = per_id.setdefault(int(row["ID"]), {"name": row["Kernel Name"]})
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        scale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9,
                 "Tbyte": 1e12}.get(u, 1.0)
        e[row["Metric Name"]] = v * scale
    ids = list(per_id)
    half = len(ids) // 2
    second = [per_id[i] for i in ids[half:]]           # the two steps launch the same sequence
    agg = collections.OrderedDict()
    for e in second:
        name = e["name"].split("(")[0].replace("<unnamed>::", "").replace("void ", "")[:70]
        a = agg.setdefault(name, {"launches": 0, "ms": 0.0, "dram_read": 0.0, "dram_write": 0.0})
        a["launches"] += 1
        a["ms"] += e.get("gpu__time_duration.sum", 0.0)
        a["dram_read"] += e.get("dram__bytes_read.sum", 0.0)
        a["dram_write"] += e.get("dram__bytes_write.sum", 0.0)
    tot_ms = sum(a["ms"] for a in agg.values())
    tot_b = sum(a["dram_read"] + a["dram_write"] for a in agg.values())
    print(f"{'ms':>9} {'share':>6} {'n':>4} {'GB read':>8} {'GB write':>8}  kernel")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
        print(f"{a['ms']:9.3f} {100 * a['ms'] / tot_ms:5.1f}% {a['launches']:4d} {a['dram_read'] / 1e9:8.3f} {a['dram_write'] / 1e9:8.3f}  {k}")
    print(f"step total: {tot_ms:.3f} ms (cold-cache, serialised), {tot_b / 1e9:.2f} GB of DRAM traffic over {len(second)} launches")
    if out:
        json.dump({"step_total_dram_bytes": tot_b, "step_total_ms_under_ncu": tot_ms, "launches": len(second),
                   "kernels": {k: {kk: (round(vv, 4) if kk == "ms" else int(vv)) for kk, vv in a.items()} for k, a in agg.items()}},
                  open(out, "w"), indent=1)

if __name__ == "__main__":
    main(*sys.argv[1:3])
