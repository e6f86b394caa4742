#!/usr/bin/env python
"""Print the key numbers of bench JSON lines: python tools/show_bench.py gpurun_out/bench_*.json"""
import json, sys
for p in sys.argv[1:]:
    try:
        d = json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e:
        print(p, "ERR", e); continue
    r = d.get("roofline", {}); e = d.get("e2e", {}); c = d.get("cpu_baseline") or {}
    print(f"{p.split('/')[-1]:24s} value {d['value']:9.1f} {d['unit']}  ms/step {d['ms_per_step']:8.3f}  e2e {e.get('value',0):8.1f}  "
          f"kern {r.get('kernel','')}: {r.get('kernel_ms',0):.3f} ms  {r.get('achieved',0):.2f}/{r.get('peak',0):.1f} TF = {100*r.get('frac',0):.1f}%  cpu {c.get('value',0):.4g}")
