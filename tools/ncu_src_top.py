#!/usr/bin/env python
"""Attribute ncu per-instruction samples to source lines.

  ncu -i X.ncu-rep --page source --csv > sass.csv
  cuobjdump -xelf all libmbots.so ; nvdisasm -g -c sim_kernels.sm_100a.cubin > sk.asm
  python tools/ncu_src_top.py sass.csv sk.asm <mangled-kernel-substring> [topN]

The ncu CSV has one row per SASS instruction in address order; nvdisasm -g has the same instructions with
`//## File "...", line N` markers (also for inlined code). They are joined by instruction index.
"""
import csv, re, sys, collections

sass_csv, asm, kern = sys.argv[1], sys.argv[2], sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 30
rows = list(csv.reader(open(sass_csv)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
cS, cSamp, cInst = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
stall_cols = [i for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
ins = [r for r in rows[hi + 1:] if len(r) >= len(h)]

# line info from nvdisasm
lines, cur, on = [], ("?", 0), False
for ln in open(asm):
    if ln.startswith("//---") and ".text." in ln:
        on = kern in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
        lines.append(cur)
if len(lines) != len(ins):
    print(f"warning: {len(lines)} disassembled instructions vs {len(ins)} profiled", file=sys.stderr)
n = min(len(lines), len(ins))
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for k in range(n):
    r = ins[k]
    a = agg[lines[k]]
    a[0] += int(r[cSamp] or 0)
    a[1] += int(r[cInst] or 0)
    for i in stall_cols:
        v = int(r[i] or 0)
        if v:
            a[2][h[i][6:]] += v
ts = sum(a[0] for a in agg.values()) or 1
ti = sum(a[1] for a in agg.values()) or 1
print(f"total samples {ts}, warp instructions {ti}")
srcs = {}
def src(f, l):
    import os, glob
    if f not in srcs:
        cand = glob.glob(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "**", f), recursive=True)
        srcs[f] = open(cand[0]).read().split("\n") if cand else []
    s = srcs[f]
    return s[l - 1].strip()[:80] if 0 < l <= len(s) else ""
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
    st = " ".join(f"{k}={v}" for k, v in a[2].most_common(3))
    print(f"{100*a[0]/ts:5.1f}% smp {100*a[1]/ti:5.1f}% ins  {f}:{l:<4} {src(f, l):80s} [{st}]")
