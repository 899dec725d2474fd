#!/bin/bash
# 200 000-voxel rows (cluster of 16, BASELINE configs[4] row length) on one GPU: 40 GB = 50 000 rows
mkdir -p gpurun_out
B="python bench.py --npixel 50000 --nvoxel 200000 --steps 2 --warmup 3 --no-cpu-baseline --sweep fused"
echo "== default =="; timeout 300 $B 2>/dev/null | python tools/bench_brief.py
echo "== per-sweep launches =="; SART_PERSISTENT=0 timeout 300 $B 2>/dev/null | python tools/bench_brief.py
echo "== 3 stages =="; SART_FUSED_STAGES=3 timeout 300 $B 2>/dev/null | python tools/bench_brief.py
echo "== V=100000 (cluster 8) =="; timeout 300 python bench.py --npixel 100000 --nvoxel 100000 --steps 2 --warmup 3 --no-cpu-baseline --sweep fused 2>/dev/null | python tools/bench_brief.py
CMD="python bench.py --npixel 50000 --nvoxel 200000 --steps 1 --warmup 3 --iterations 2 --no-cpu-baseline --sweep fused"
SART_PERSISTENT=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 4 -c 1 -f -o gpurun_out/prof_c16 env SART_PERSISTENT=0 $CMD > gpurun_out/ncu_c16.log 2>&1
tail -2 gpurun_out/ncu_c16.log
