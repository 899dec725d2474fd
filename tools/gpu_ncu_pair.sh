#!/bin/bash
# Full ncu capture of the two-frame fused sweep (sart_solve_batch) on the bench matrix.
mkdir -p gpurun_out
CMD="python tools/run_config.py --config c2 --batch 2 --iterations 3"
$CMD > gpurun_out/plain_ncu_pair.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fused_sweep -s 1 -c 1 -f -o gpurun_out/prof_pair $CMD > gpurun_out/ncu_pair.log 2>&1
tail -3 gpurun_out/ncu_pair.log
ls -la gpurun_out/prof_pair.ncu-rep
