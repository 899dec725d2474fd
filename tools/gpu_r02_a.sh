#!/bin/bash
# round 2, first single-GPU call: every GPU test, smoke plain and under ncu (launch list), the default bench line with its
# check block, the run-ahead producer on/off, cluster-size tables for the wide rows
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit,memory.total --format=csv > gpurun_out/r02_box.txt; nproc >> gpurun_out/r02_box.txt; free -g | head -2 >> gpurun_out/r02_box.txt
echo "== pytest -m gpu =="
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r02a_pytest_gpu.txt
echo "== smoke plain, then under ncu (launch list) =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02a_smoke.txt 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r02a_smoke_launches.csv python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02a_smoke_ncu.txt 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r02a_smoke.txt; tail -3 gpurun_out/r02a_smoke_ncu.txt; grep -c fused_solve gpurun_out/r02a_smoke_launches.csv
env | grep -i -E "inject|nsight|profiler" | head
echo "== bench (default) =="
timeout 900 python bench.py > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "rc=$?"; python tools/bench_brief.py < gpurun_out/r02a_bench.json; tail -3 gpurun_out/r02a_bench.err
echo "== run-ahead producer off =="
SART_PREFETCH=0 timeout 600 python bench.py --no-cpu-baseline --no-check > gpurun_out/r02a_bench_noprefetch.json 2>/dev/null; python tools/bench_brief.py < gpurun_out/r02a_bench_noprefetch.json
echo "== c1 run-ahead on / off =="
timeout 300 python bench.py --workload c1 --no-cpu-baseline --no-check 2>/dev/null | python tools/bench_brief.py
SART_PREFETCH=0 timeout 300 python bench.py --workload c1 --no-cpu-baseline --no-check 2>/dev/null | python tools/bench_brief.py
echo "== cluster tables =="
timeout 600 python tools/cluster_table.py --nvoxel 50000 --gb 20 --clusters 4,5,6,7 2>&1 | grep CLUSTER | tee gpurun_out/r02a_clusters_50k.txt
timeout 600 python tools/cluster_table.py --nvoxel 100000 --gb 20 --clusters 7,8,9,10,12 2>&1 | grep CLUSTER | tee gpurun_out/r02a_clusters_100k.txt
timeout 600 python tools/cluster_table.py --nvoxel 200000 --gb 20 --clusters 14,15,16 2>&1 | grep CLUSTER | tee gpurun_out/r02a_clusters_200k.txt
