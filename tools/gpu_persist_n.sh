#!/bin/bash
# multi-GPU check of the whole-solve kernel: tests, then the bench with and without it
N=${1:-2}
echo "== multi-GPU tests =="
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -4
for p in 1 0; do
  echo "== bench N=$N SART_PERSISTENT=$p =="
  SART_PERSISTENT=$p timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 2>/dev/null | python tools/bench_brief.py
done
