#!/bin/bash
mkdir -p gpurun_out
echo "== fused tests =="
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "fused" 2>&1 | tail -5
SART_FUSED_THREADS=224 timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "fused" 2>&1 | tail -5
echo "== bench fused 480 =="
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep fused 2>&1 | python tools/bench_brief.py
echo "== bench fused 224 =="
SART_FUSED_THREADS=224 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep fused 2>&1 | python tools/bench_brief.py
for st in 3; do echo "== fused 480 stages=$st =="; SART_FUSED_STAGES=$st timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep fused 2>&1 | python tools/bench_brief.py; done
