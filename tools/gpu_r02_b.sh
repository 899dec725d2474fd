#!/bin/bash
# where the time between two sweeps goes on one GPU (diagnostic timers of thread 0 of CTA 0)
for W in c1 c2; do
  echo "== $W =="
  SART_TAIL_DEBUG=1 timeout 600 python bench.py --workload $W --no-cpu-baseline --no-check --steps 3 2> gpurun_out/r02b_$W.err | python tools/bench_brief.py
  grep "tail debug" gpurun_out/r02b_$W.err | tail -2
done
timeout 600 python -m pytest tests/test_host_cpp_gpu.py tests/test_parity_gpu.py -m gpu -x -q -k "template or whole_solve or golden" 2>&1 | tail -3
