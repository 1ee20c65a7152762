"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg, tot = collections.OrderedDict(), 0.0
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        v *= {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}[row["Metric Unit"]]
        a = agg.setdefault(row["Kernel Name"][:90], [0, 0.0])
        a[0] += 1
        a[1] += v
        tot += v
    print(f"{'ms':>10} {'n':>5} {'share':>6}  kernel")
    for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
        print(f"{v / 1e6:10.3f} {n:5d} {100 * v / tot:5.1f}%  {k}")
    print(f"total {tot / 1e6:.3f} ms over {sum(a[0] for a in agg.values())} launches")


if __name__ == "__main__":
    main(sys.argv[1])
