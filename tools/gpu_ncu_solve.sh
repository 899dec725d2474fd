#!/bin/bash
# ncu launch list + one full capture of the whole-solve kernel. Nsight Compute cannot replay the cooperative cluster
# launch, so these two passes run the identical kernel and grid without the cooperative attribute (SART_COOPERATIVE=0).
mkdir -p gpurun_out
export SART_COOPERATIVE=0
CMD="python bench.py --steps 1 --warmup 3 --iterations 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_c.log 2>&1 || exit 1
python tools/bench_brief.py < gpurun_out/plain_c.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_solve.csv $CMD > gpurun_out/ncu_c.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_solve.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_solve -s 3 -c 1 -f -o gpurun_out/prof_solve $CMD > gpurun_out/ncu_full_c.log 2>&1
tail -3 gpurun_out/ncu_full_c.log
ls -la gpurun_out/prof_solve.ncu-rep
