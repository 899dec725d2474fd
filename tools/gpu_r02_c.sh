#!/bin/bash
# after the statistics / tail changes: every single-GPU test, the tail diagnostics, then (with 2 GPUs) the multi-GPU tests
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/r02c_pytest_gpu.txt
for W in c1 c2; do
  SART_TAIL_DEBUG=1 timeout 600 python bench.py --workload $W --no-cpu-baseline --no-check --steps 3 2> gpurun_out/r02c_$W.err | python tools/bench_brief.py
  grep "tail debug" gpurun_out/r02c_$W.err | tail -1
done
