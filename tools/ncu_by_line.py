"""Aggregates an `ncu --page source --csv` export (SASS rows) by CUDA source line, using the line table
nvdisasm prints for the same cubin:
    cuobjdump -xelf all kmi_ops_k31.o; nvdisasm --print-line-info X.cubin > dis.txt
    ncu -i rep --page source --csv --kernel-name regex:NAME > sass.csv
    python tools/ncu_by_line.py sass.csv dis.txt MANGLED_SUBSTRING [top]
Prints, per file:line, warp instructions executed and stall samples (share of the kernel)."""
import csv
import re
import sys
from collections import defaultdict

sass_csv, dis, mangled = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
addr2line = {}
cur = None
infn = False
for ln in open(dis):
    if ln.startswith("//--------------------- .text."):
        infn = mangled in ln
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/", ln)
    if m:
        addr2line[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(sass_csv)))
hdr = None
first = None
agg = defaultdict(lambda: [0, 0, 0])
for r in rows:
    if r and r[0] == "Address":
        if hdr is not None:
            break  # first kernel instance only
        hdr = r
        ia, ii, isamp, ith = r.index("Address"), r.index("Instructions Executed"), r.index("# Samples"), r.index("Thread Instructions Executed")
        continue
    if hdr is None or len(r) < len(hdr):
        continue
    a = int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia])
    if first is None:
        first = a
    k = addr2line.get(a - first, ("?", 0))
    agg[k][0] += int(r[ii] or 0)
    agg[k][1] += int(r[isamp] or 0)
    agg[k][2] += int(r[ith] or 0)
ti = sum(v[0] for v in agg.values())
ts = sum(v[1] for v in agg.values())
print("total warp instr %.3e  samples %d" % (ti, ts))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-18s:%4d  instr %5.1f%%  samples %5.1f%%  thr/warp %4.1f" % (k[0], k[1], 100.0 * v[0] / ti, 100.0 * v[1] / max(ts, 1), v[2] / max(v[0], 1)))
if len(sys.argv) > 5:   # region summary: "file:lo-hi=name,..."
    print("-- regions")
    for spec in sys.argv[5].split(","):
        rng, name = spec.split("=")
        f, lh = rng.split(":")
        lo, hi = map(int, lh.split("-"))
        si = sum(v[0] for k, v in agg.items() if k[0].startswith(f) and lo <= k[1] <= hi)
        ss = sum(v[1] for k, v in agg.items() if k[0].startswith(f) and lo <= k[1] <= hi)
        print("%-28s instr %5.1f%%  samples %5.1f%%" % (name, 100.0 * si / ti, 100.0 * ss / max(ts, 1)))
