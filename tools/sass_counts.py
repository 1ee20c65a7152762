"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native kernel (B200_PROFILING.md):
UTC*MMA (tcgen05.mma), LDTM/STTM (tcgen05.ld/st), UTCCP (tcgen05.cp), UTMALDG/UTMASTG (TMA), HMMA (legacy mma.sync).
Usage: python tools/sass_counts.py [path/to/lib.so] > profiles/r02_sass_counts.txt   (runs without a GPU)"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "rnnt-transducer_b200", "librnnt_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
pats = ["UTCHMMA.2CTA", "UTCHMMA", "UTCCP", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTCBAR", "SYNCS", "HMMA", "FFMA2", "MUFU.EX2", "REDG", "ATOMG"]
rows, cur, cnt = [], None, None
for ln in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        if cur: rows.append((cur, cnt))
        cur, cnt = m.group(1), {p: 0 for p in pats}
        continue
    if cur is None: continue
    ins = re.search(r"/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]*)", ln)
    if not ins: continue
    op = ins.group(1)
    for p in pats:
        if p == "UTCHMMA":
            if op.startswith("UTCHMMA") and ".2CTA" not in op: cnt[p] += 1
        elif p == "UTCHMMA.2CTA":
            if op.startswith("UTCHMMA") and ".2CTA" in op: cnt[p] += 1
        elif op.startswith(p): cnt[p] += 1
if cur: rows.append((cur, cnt))
demangle = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True, text=True).stdout.splitlines()
print(f"# {os.path.relpath(so, ROOT)}: cuobjdump -sass, instruction counts per kernel (sm_100a)")
for (name, cnt), dm in zip(rows, demangle):
    short = re.sub(r"\(anonymous namespace\)::", "", dm).split("(")[0]
    nz = {k: v for k, v in cnt.items() if v}
    print(f"{short}: " + (", ".join(f"{k}={v}" for k, v in nz.items()) or "-"))
