#!/bin/bash
# end-of-round check on one GPU: every GPU test, smoke, the default bench line, the 150 GB shard of configs[4] alone
mkdir -p gpurun_out
echo "== pytest -m gpu =="
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_final.txt
echo "== smoke =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 | tee gpurun_out/smoke_final.txt
echo "== bench (default) =="
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; python tools/bench_brief.py < gpurun_out/bench_final.json
echo "== configs[4] shard (187 500 x 200 000 = 150 GB) on one GPU =="
timeout 900 python tools/run_config.py --config c5 --scale 0.125 --iterations 5 2>&1 | grep CONFIG | tee gpurun_out/c5_shard_1gpu.json | cut -c1-1200
