#!/usr/bin/env python
"""print the per-stage table of one or more bench.py JSON lines"""
import json, sys
for p in sys.argv[1:]:
    try:
        d = json.loads(open(p).read().strip().splitlines()[-1])
    except Exception as e:
        print(p, "unreadable:", e); continue
    print(f"== {p}: {d['value']/1e9:.2f} G/s  {d['ms_per_step']:.3f} ms/step  e2e {d['e2e']['value']/1e9:.3f} G/s {d['e2e']['ms_per_step']:.1f} ms  frac {d['roofline']['frac']:.3f} ({d['roofline']['kernel']})")
    for k, v in d['roofline']['kernels'].items():
        print(f"   {k:45s} {v['ms_per_step']:.3f} ms/step share {v['share']:.3f} {v['GBps']:.0f} GB/s")
