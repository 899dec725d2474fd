#!/bin/bash
mkdir -p gpurun_out
echo "== pytest -m gpu =="
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/pytest_gpu.txt
echo "== smoke =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== bench (auto) =="
timeout 600 python bench.py > gpurun_out/bench_auto.json 2> gpurun_out/bench_auto.err; python tools/bench_brief.py < gpurun_out/bench_auto.json; tail -3 gpurun_out/bench_auto.err
echo "== ncu launch list =="
CMD="python bench.py --steps 1 --warmup 3 --iterations 3 --no-cpu-baseline"
$CMD > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_fused.csv $CMD > gpurun_out/ncu.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_fused.csv
echo "== ncu full (fused kernel) =="
bash tools/gpu_ncu_fused.sh fused fused_sweep
