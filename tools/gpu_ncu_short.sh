#!/bin/bash
# Full ncu capture of the fused sweep on a short-row shape (default 4096 voxels, ~2 GB matrix).
mkdir -p gpurun_out
V=${1:-4096}
P=$((2000000000 / V))
CMD="python bench.py --npixel $P --nvoxel $V --steps 1 --warmup 3 --iterations 2 --no-cpu-baseline --sweep fused"
$CMD > gpurun_out/plain_ncu_short.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 4 -c 1 -f -o gpurun_out/prof_short_$V $CMD > gpurun_out/ncu_short.log 2>&1
tail -3 gpurun_out/ncu_short.log
ls -la gpurun_out/prof_short_$V.ncu-rep
