#!/bin/bash
# round 2, single GPU: all GPU tests, the series (configs[3]) on one GPU, ncu captures of the whole-solve kernel on the
# 40 GB shard (cluster of 4) and on 200 000-voxel rows (cluster of 16), the race against the reference's own GPU solver
mkdir -p gpurun_out
echo "== pytest -m gpu =="
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 | tee gpurun_out/r02d_pytest_gpu.txt
echo "== bench default =="
timeout 900 python bench.py > gpurun_out/r02d_bench.json 2> gpurun_out/r02d_bench.err; echo "rc=$?"; python tools/bench_brief.py < gpurun_out/r02d_bench.json | cut -c1-1500
echo "== c1 =="
timeout 600 python bench.py --workload c1 > gpurun_out/r02d_bench_c1.json 2>/dev/null; python tools/bench_brief.py < gpurun_out/r02d_bench_c1.json | cut -c1-1200
echo "== c4 series, 1000 frames, one GPU =="
timeout 900 python bench.py --workload c4 --frames 1000 > gpurun_out/r02_c4_series_n1.json 2> gpurun_out/r02_c4_series_n1.err; echo "rc=$?"; python tools/bench_brief.py < gpurun_out/r02_c4_series_n1.json | cut -c1-1500
echo "== ncu: whole-solve kernel, 40 GB shard =="
CMD="python bench.py --iterations 2 --steps 1 --warmup 3 --no-check --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/r02d_plain_c2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_solve -s 3 -c 1 -f -o gpurun_out/r02_prof_solve_c2 $CMD > gpurun_out/r02d_ncu_c2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r02d_ncu_c2.log
echo "== ncu: whole-solve kernel, 200 000-voxel rows (cluster of 16) =="
CMD="python bench.py --npixel 25000 --nvoxel 200000 --iterations 2 --steps 1 --warmup 3 --no-check --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/r02d_plain_c16.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:fused_solve -s 3 -c 1 -f -o gpurun_out/r02_prof_solve_c16 $CMD > gpurun_out/r02d_ncu_c16.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/r02d_ncu_c16.log
echo "== race against the reference's GPU solver: configs[0] =="
timeout 600 python tools/race_reference_gpu.py --rows 8192 --nvoxel 4096 --iterations 100 2>&1 | grep -E "RACE|rror" | tee gpurun_out/r02_reference_gpu_race_c1.json | cut -c1-1500
echo "== race: the full 40 GB configs[1] matrix =="
timeout 1500 python tools/race_reference_gpu.py --rows 200000 --nvoxel 50000 --iterations 10 > gpurun_out/r02_race_c2.log 2>&1; echo "rc=$?"
grep -E "RACE" gpurun_out/r02_race_c2.log | tee gpurun_out/r02_reference_gpu_race_c2.json | cut -c1-1500; tail -5 gpurun_out/r02_race_c2.log | cut -c1-400
