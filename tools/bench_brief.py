#!/usr/bin/env python
"""Prints the few numbers of a bench.py JSON line that matter when comparing runs."""
import json
import sys

for line in sys.stdin:
    if not line.startswith("{"):
        if line.strip():
            print(line.rstrip()[:300])
        continue
    d = json.loads(line)
    k = d["roofline"]["kernels"]
    parts = [f"N={d['n_gpus']} sweep={d['config'].get('sweep')} exchange={d['config'].get('exchange')} value={d['value']:.2f} it/s e2e={d['e2e']['value']:.2f} ms/step={d['ms_per_step']:.3f}"]
    for name, v in k.items():
        if v["launches"]:
            parts.append(f"{name}: {v['ms']:.3f} ms x{v['launches']} alg {v['gbs']:.0f} GB/s hbm {v['hbm_gbs']:.0f} GB/s share {v['share_of_step']:.3f}")
    parts.append(f"between_sweeps={d.get('between_sweeps_us', 0):.1f} us")
    parts.append(f"clocks={d.get('clocks')}")
    print(" | ".join(parts))
