#!/usr/bin/env python
"""Prints the few numbers of a bench.py JSON line that matter when comparing runs: bench_brief.py FILE... (or on stdin)."""
import fileinput
import json

for line in fileinput.input():
    if not line.startswith("{"):
        if line.strip():
            print(line.rstrip()[:300])
        continue
    d = json.loads(line)
    parts = [f"N={d['n_gpus']} sweep={d['config'].get('sweep')} exchange={d['config'].get('exchange')} value={d['value']:.2f} it/s "
             f"e2e={d['e2e']['value']:.2f} ms/step={d['ms_per_step']:.3f}"]
    for name, v in d.get("roofline", {}).get("kernels", {}).items():
        if v["launches"]:
            parts.append(f"{name}: {v['ms']:.3f} ms x{v['launches']} hbm {v['hbm_gbs']:.0f} GB/s alg {v['algorithmic_gbs']:.0f} GB/s "
                         f"share {v['share_of_step']:.3f}")
    parts.append(f"between_sweeps={d.get('between_sweeps_us', 0):.1f} us {d.get('between_sweeps_phases_us')}")
    parts.append(f"geometry={d.get('geometry')}")
    if d.get("series"):
        parts.append(f"series={d['series']}")
    parts.append(f"check={d.get('check')}")
    if d.get("extra"):
        parts.append(f"extra={d['extra']}")
    parts.append(f"clocks={d.get('clocks')}")
    print(" | ".join(parts))
