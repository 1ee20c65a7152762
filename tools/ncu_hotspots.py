"""Per-kernel hot spots from `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda`:
aggregates warp-stall samples by CUDA source line and lists the top SASS instructions."""
import csv
import sys
import collections


def main(path, kernel_filter="", top=25):
    rows = list(csv.reader(open(path)))
    i = 0
    blocks = []
    while i < len(rows):
        if rows[i] and rows[i][0] == "File Path":
            fpath = rows[i][1]
            fn = rows[i + 1][1]
            hdr = rows[i + 2]
            j = i + 3
            body = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "File Path"):
                body.append(rows[j]); j += 1
            blocks.append((fpath, fn, hdr, body))
            i = j
        else:
            i += 1
    for fpath, fn, hdr, body in blocks:
        if kernel_filter not in fn:
            continue
        if "Address" not in hdr:
            continue
        a_i = hdr.index("Address")
        s_i = hdr.index("# Samples")
        src_cols = [k for k, h in enumerate(hdr) if h == "Source"]
        stall_cols = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = 0
        per = []
        for r in body:
            if len(r) <= s_i:
                continue
            try:
                n = int(r[s_i])
            except ValueError:
                continue
            tot += n
            stalls = {hdr[k]: int(r[k]) for k in stall_cols if r[k].isdigit() and int(r[k]) > 0}
            per.append((n, r[src_cols[-1]] if len(src_cols) > 1 else "", r[src_cols[0]], r[0], stalls))
        print("=" * 100)
        print(fn[:110], "| file", fpath.split("/")[-1], "| samples", tot)
        per.sort(key=lambda x: -x[0])
        for n, sass, src, line, stalls in per[:top]:
            top_st = sorted(stalls.items(), key=lambda kv: -kv[1])[:3]
            print(f"{n:7d} {100.0 * n / max(tot, 1):5.1f}%  L{line:>4} {sass[:70]:70s} {top_st}")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "", int(sys.argv[3]) if len(sys.argv) > 3 else 25)
