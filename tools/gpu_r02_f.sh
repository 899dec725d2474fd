#!/bin/bash
# regression after the batch changes (log solver, folded per-frame launches): batch timings on the 40 GB matrix and on
# configs[0]'s size, every GPU test, smoke, the default bench line
mkdir -p gpurun_out
timeout 600 python tools/batch_bench.py 2>&1 | grep -E "BATCH|rror" | tee gpurun_out/r02f_batch_c2.json | cut -c1-900
timeout 300 python tools/batch_bench.py --npixel 8192 --nvoxel 4096 --frames 64 --iterations 50 2>&1 | grep -E "BATCH|rror" | tee gpurun_out/r02f_batch_c1.json | cut -c1-900
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02f_pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/r02f_bench.json 2> gpurun_out/r02f_bench.err; echo "bench rc=$?"
timeout 20 python tools/bench_brief.py gpurun_out/r02f_bench.json 2>&1 | cut -c1-1500
