#!/bin/bash
# four vs eight frames per matrix pass (sart_solve_batch) on the 40 GB matrix, and the batch tests
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "batch" 2>&1 | tail -3
for G in 8 4; do
  SART_BATCH_FRAMES=$G timeout 600 python tools/batch_bench.py 2>&1 | grep -E "BATCH|rror" | tee gpurun_out/r02_batch_c2_group$G.json | cut -c1-1200
done
