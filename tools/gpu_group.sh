#!/bin/bash
echo "== batch tests =="
timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "batch_of" 2>&1 | tail -3
echo "== batch on c2 =="
timeout 300 python tools/run_config.py --config c2 --batch 8 --iterations 10 2>&1 | grep CONFIG | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l[7:]); print({k: d[k] for k in ('speedup','group_iteration_ms','group_fp_ms','group_bp_ms','single_sweep_ms','frame_iterations_per_s_batch','frame_iterations_per_s_one_by_one','max_rel_l2_batch_vs_single')})"
