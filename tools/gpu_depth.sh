#!/bin/bash
echo "== fused tests =="
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "fused or batch_of or randomised" 2>&1 | tail -3
echo "== batch on c2 =="
timeout 300 python tools/run_config.py --config c2 --batch 8 --iterations 10 2>&1 | grep CONFIG | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l[7:]); print({k: d[k] for k in ('speedup','group_iteration_ms','group_fp_ms','group_bp_ms','single_sweep_ms','frame_iterations_per_s_batch','frame_iterations_per_s_one_by_one','max_rel_l2_batch_vs_single')})"
echo "== short rows: depth 1 vs 3 =="
for V in 4096 8192 16384; do
  P=$((2000000000 / V))
  for dp in 1 3; do
    SART_FUSED_DEPTH=$dp timeout 200 python bench.py --npixel $P --nvoxel $V --steps 3 --warmup 3 --no-cpu-baseline --sweep fused 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('V=$V depth=$dp it/s', round(d['value'],1))"
  done
done
echo "== c2 depth 3 forced (4 stages) =="
SART_FUSED_DEPTH=3 timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python tools/bench_brief.py
