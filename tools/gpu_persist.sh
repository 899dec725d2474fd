#!/bin/bash
echo "== whole-solve kernel tests =="
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "whole_solve or fixed_iterations or golden or convergence" 2>&1 | tail -5
echo "== bench (default) =="
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py
echo "== bench per-sweep launches =="
SART_PERSISTENT=0 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py
echo "== c1 =="
timeout 300 python bench.py --workload c1 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py
SART_PERSISTENT=0 timeout 300 python bench.py --workload c1 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py
