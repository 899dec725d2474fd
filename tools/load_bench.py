#!/usr/bin/env python
"""Throughput of the file -> GPU shard path (sart_load_rows_file: pread into three pinned staging buffers overlapped
with asynchronous copies) against a host-resident upload (sart_upload_rows), on a scratch file of --gb gigabytes."""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=4.0)
    ap.add_argument("--nvoxel", type=int, default=50000)
    args = ap.parse_args()
    sart = load_package()
    V = args.nvoxel
    P = int(args.gb * 1e9 / (4 * V))
    rng = np.random.default_rng(0)
    H = rng.random((P, V), dtype=np.float32)
    out = {"rows": P, "nvoxel": V, "gb": P * V * 4 / 1e9}
    with tempfile.TemporaryDirectory(dir=os.environ.get("TMPDIR", "/tmp")) as d:
        path = os.path.join(d, "block.bin")
        H.tofile(path)
        with sart.Engine(P, V, device=0) as eng:
            for label, mb in (("staging_64MB", "64"), ("staging_8MB", "8"), ("staging_256MB", "256")):
                os.environ["SART_STAGING_MB"] = mb
                t = time.perf_counter()
                eng.load_rows_file(path, 0, V, 0, P, 0, V)
                out[label + "_gb_s"] = out["gb"] / (time.perf_counter() - t)
            ok = bool(np.array_equal(eng.download_rows(0, 64), H[:64]) and np.array_equal(eng.download_rows(P - 64, 64), H[-64:]))
            t = time.perf_counter()
            eng.upload_rows(H)
            out["upload_rows_pageable_gb_s"] = out["gb"] / (time.perf_counter() - t)
        out["verified"] = ok
    print("LOAD " + json.dumps(out))


if __name__ == "__main__":
    main()
