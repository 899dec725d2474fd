#!/bin/bash
mkdir -p gpurun_out
echo "== fused tests (small first) =="
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "fused and not full_size and not agree" 2>&1 | tail -15
echo "== all gpu tests =="
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
echo "== bench split =="
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep split 2>&1 | python tools/bench_brief.py
echo "== bench fused =="
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep fused > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err; python tools/bench_brief.py < gpurun_out/bench_fused.json; tail -3 gpurun_out/bench_fused.err
for st in 2 3; do echo "== fused stages=$st =="; SART_FUSED_STAGES=$st timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep fused 2>&1 | python tools/bench_brief.py; done
echo "== fused cluster=8 =="; SART_FUSED_CLUSTER=8 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --sweep fused 2>&1 | python tools/bench_brief.py
