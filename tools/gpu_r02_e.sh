#!/bin/bash
# sart_solve_batch: four vs eight frames per matrix pass on the 40 GB matrix, ncu capture of the two eight-frame
# projection kernels, then every GPU test
mkdir -p gpurun_out
for G in 8 4; do
  SART_BATCH_FRAMES=$G timeout 600 python tools/batch_bench.py 2>&1 | grep -E "BATCH|rror" | tee gpurun_out/r02_batch_c2_group$G.json | cut -c1-1200
done
CMD="python tools/batch_bench.py --frames 8 --iterations 2"
timeout 300 $CMD > gpurun_out/r02e_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:batch_ -s 4 -c 2 -f -o gpurun_out/r02_prof_batch8 $CMD > gpurun_out/r02e_ncu.log 2>&1
echo "ncu rc=$?"; tail -n 2 gpurun_out/r02e_ncu.log | cut -c1-300
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02e_pytest_gpu.txt
