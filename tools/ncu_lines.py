#!/usr/bin/env python
"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump by CUDA source line.
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv; python tools_ncu_lines.py src.csv [top]"""
import csv, sys, collections
path = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
# find header rows
agg = collections.OrderedDict(); tot_inst = 0; tot_samp = 0
cur_file = None; hdr = None
opc = collections.Counter()
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur_file = r[1].split("/")[-1]; continue
    if r[0] == "Function Name": continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None: continue
    d = dict(zip(hdr, r))
    # columns are duplicated ("Source" twice): use indices
    try:
        line = r[0]; src = r[1]; sass = r[3]
        samples = int(r[6] or 0); inst = int(r[7] or 0)
    except (ValueError, IndexError):
        continue
    if line:                       # CUDA source row: counts aggregated over its SASS
        key = (cur_file, line, src.strip()[:90])
        a = agg.setdefault(key, [0, 0])
        a[0] += inst; a[1] += samples
        tot_inst += inst; tot_samp += samples
    elif sass and sass != "-":     # SASS row (some rows carry no executed count in the CSV)
        t = sass.split()
        op = t[1] if t[0].startswith("@") else t[0]
        opc[op.split(".")[0]] += samples
print("total warp-instructions", tot_inst, "samples", tot_samp)
for (f, l, s), (i, sm) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100*i/max(tot_inst,1):5.1f}% inst {100*sm/max(tot_samp,1):5.1f}% samp  {f}:{l}  {s}")
print("--- opcode mix")
for op, i in opc.most_common(30):
    print(f"{100*i/max(sum(opc.values()),1):5.1f}% of samples  {op}")
