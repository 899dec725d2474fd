#!/bin/bash
# quick regression: all single-GPU tests, the batch measurement, the default bench line
mkdir -p gpurun_out
echo "== pytest -m gpu =="
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_d.txt
echo "== batch on c2 =="
timeout 300 python tools/run_config.py --config c2 --batch 8 --iterations 10 2>&1 | grep CONFIG | tee gpurun_out/c2_batch.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l[7:]); print({k: d[k] for k in ('speedup','group_iteration_ms','group_fp_ms','group_bp_ms','single_sweep_ms','frame_iterations_per_s_batch','frame_iterations_per_s_one_by_one','max_rel_l2_batch_vs_single')})"
echo "== bench =="
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py
