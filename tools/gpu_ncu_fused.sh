#!/bin/bash
# One full ncu capture of the fused sweep kernel (and of the two split kernels for comparison), on the bench workload.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --iterations 2 --no-cpu-baseline --sweep ${1:-fused}"
$CMD > gpurun_out/plain_ncu.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"${2:-fused_sweep}" -s 4 -c 1 -f -o gpurun_out/prof_${1:-fused} $CMD > gpurun_out/ncu_full.log 2>&1
tail -5 gpurun_out/ncu_full.log
ls -la gpurun_out/*.ncu-rep
