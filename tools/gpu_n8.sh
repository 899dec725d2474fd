#!/bin/bash
# 8-GPU measurements: bench with / without the whole-solve kernel, time series (c4), 400 GB matrix (c3)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
for p in 1 0; do
  echo "== bench N=8 SART_PERSISTENT=$p =="
  SART_PERSISTENT=$p timeout 300 $TR 29521 bench.py --gpus 8 --steps 5 --warmup 3 2>/dev/null | tee gpurun_out/bench_n8_persist$p.json | python tools/bench_brief.py
done
echo "== c4 N=8 =="
timeout 300 $TR 29522 tools/run_config.py --config c4 --frames 100 2>/dev/null | grep CONFIG | tee gpurun_out/c4_n8.json | cut -c1-900
echo "== c3 N=8 =="
timeout 400 $TR 29523 tools/run_config.py --config c3 --iterations 5 2>/dev/null | grep CONFIG | tee gpurun_out/c3_n8.json | cut -c1-900
