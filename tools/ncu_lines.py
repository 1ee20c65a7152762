"""Top CUDA source lines by warp-stall samples, per kernel, from
`ncu -i X.ncu-rep --page source --csv --print-source sass,cuda > X.csv`."""
import csv, sys, collections

def main(path, top=30):
    rows = list(csv.reader(open(path)))
    i, per_kernel = 0, collections.OrderedDict()
    while i < len(rows):
        if rows[i] and rows[i][0] == "File Path":
            f, fn, hdr = rows[i][1].split("/")[-1], rows[i + 1][1], rows[i + 2]
            j = i + 3
            s_i = hdr.index("# Samples"); ie_i = hdr.index("Instructions Executed")
            stall = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
            while j < len(rows) and not (rows[j] and rows[j][0] == "File Path"):
                r = rows[j]; j += 1
                if len(r) <= s_i or r[2] != "-" or not r[s_i].isdigit():
                    continue
                st = sorted(((int(r[k]), hdr[k][6:]) for k in stall if r[k].isdigit() and int(r[k])), reverse=True)[:3]
                per_kernel.setdefault(fn, []).append((int(r[s_i]), f, r[0], r[1].strip()[:90], r[ie_i], st))
            i = j
        else:
            i += 1
    for fn, lines in per_kernel.items():
        tot = sum(l[0] for l in lines)
        print("=" * 120); print(fn[:100], "| samples", tot)
        for n, f, ln, src, ie, st in sorted(lines, reverse=True)[:top]:
            print(f"{n:7d} {100.0*n/tot:5.1f}% {f[:14]:14s}:{ln:>4} inst={ie:>9} {src:90s} {st}")

if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30)
