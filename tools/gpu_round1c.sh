#!/bin/bash
# Round-1 final evidence on one GPU: all GPU tests, smoke, default bench line, ncu launch list and one full capture of
# the whole-solve kernel (each ncu pass only after the same command has exited 0 without ncu).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/box_c.txt; nproc >> gpurun_out/box_c.txt
echo "== pytest -m gpu =="
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -6 | tee gpurun_out/pytest_gpu_c.txt
echo "== smoke =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== bench (default) =="
timeout 600 python bench.py > gpurun_out/bench_solve.json 2> gpurun_out/bench_solve.err; python tools/bench_brief.py < gpurun_out/bench_solve.json; tail -2 gpurun_out/bench_solve.err
echo "== bench --impl reference =="
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 2>/dev/null | tee gpurun_out/bench_reference_c.json | cut -c1-400
echo "== ncu launch list =="
CMD="python bench.py --steps 1 --warmup 3 --iterations 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_c.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_solve.csv $CMD > gpurun_out/ncu_c.log 2>&1
python tools/summarize_launches.py gpurun_out/launches_solve.csv
echo "== ncu full (whole-solve kernel) =="
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_solve -s 3 -c 1 -f -o gpurun_out/prof_solve $CMD > gpurun_out/ncu_full_c.log 2>&1
tail -3 gpurun_out/ncu_full_c.log
ls -la gpurun_out/prof_solve.ncu-rep
