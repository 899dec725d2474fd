#!/bin/bash
# round 2, N GPUs of one box (gpurun --gpus N): the multi-GPU parity tests with their log kept, the default bench line with
# its check block (+ the 50 GB-per-GPU configs[2] record), the warm-started series (configs[3]) and, with 8 GPUs, configs[2]
# and configs[4] as bench workloads
N=${1:-2}
FRAMES=${2:-1000}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/r02_box_n$N.txt; nproc >> gpurun_out/r02_box_n$N.txt
echo "== multi-GPU tests =="
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -v 2>&1 | tail -12 | tee gpurun_out/r02_multi_gpu_pytest_n$N.txt
echo "== bench N=$N =="
timeout 900 $TR --master-port 29541 bench.py --gpus $N --extra-c3 on > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "rc=$?"
python tools/bench_brief.py < gpurun_out/r02_bench_n$N.json; tail -2 gpurun_out/r02_bench_n$N.err
if [ "${3:-}" != "fast" ]; then
  echo "== bench N=$N, run-ahead producer off =="
  SART_PREFETCH=0 timeout 600 $TR --master-port 29542 bench.py --gpus $N --no-check --extra-c3 off 2>/dev/null | tee gpurun_out/r02_bench_n${N}_noprefetch.json | python tools/bench_brief.py
fi
echo "== c4 series N=$N ($FRAMES frames) =="
timeout 1500 $TR --master-port 29543 bench.py --gpus $N --workload c4 --frames $FRAMES > gpurun_out/r02_c4_series_n$N.json 2> gpurun_out/r02_c4_series_n$N.err; echo "rc=$?"
python tools/bench_brief.py < gpurun_out/r02_c4_series_n$N.json; tail -2 gpurun_out/r02_c4_series_n$N.err
if [ "$N" = "8" ]; then
  echo "== c3 (400 GB) =="
  timeout 900 $TR --master-port 29544 bench.py --gpus 8 --workload c3 --extra-c3 off > gpurun_out/r02_bench_c3_n8.json 2> gpurun_out/r02_bench_c3_n8.err; echo "rc=$?"
  python tools/bench_brief.py < gpurun_out/r02_bench_c3_n8.json; tail -2 gpurun_out/r02_bench_c3_n8.err
  echo "== c5 (1.2 TB) =="
  timeout 1200 $TR --master-port 29545 bench.py --gpus 8 --workload c5 --extra-c3 off --steps 3 > gpurun_out/r02_bench_c5_n8.json 2> gpurun_out/r02_bench_c5_n8.err; echo "rc=$?"
  python tools/bench_brief.py < gpurun_out/r02_bench_c5_n8.json; tail -2 gpurun_out/r02_bench_c5_n8.err
fi
