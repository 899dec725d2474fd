#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
echo "== multi-GPU tests =="
timeout 700 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -6
echo "== c4 series N=$N =="
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 tools/run_config.py --config c4 --frames 100 --series 2>/dev/null | grep CONFIG | tee gpurun_out/c4_series_n$N.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l[7:]); print({k: d.get(k) for k in ('n_gpus','frames_per_s','iterations_total','wall_s','device_ms_total','all_converged','matrix_passes','frames_handed_their_first_iteration','sweeps_repeated','ranks_agree')})"
