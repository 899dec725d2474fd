#!/bin/bash
echo "== series tests =="
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "series or whole_solve" 2>&1 | tail -15
echo "== c4, 60 frames, 1 GPU: resident solve per frame vs series call =="
timeout 300 python tools/run_config.py --config c4 --frames 60 2>/dev/null | grep CONFIG | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l[7:]); print({k: d.get(k) for k in ('frames_per_s','iterations_total','wall_s','device_ms_total','all_converged','last_frame_rel_l2_vs_truth')})"
timeout 300 python tools/run_config.py --config c4 --frames 60 --series 2>/dev/null | grep CONFIG | tee gpurun_out/c4_series_1gpu.json | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l[7:]); print({k: d.get(k) for k in ('frames_per_s','iterations_total','wall_s','device_ms_total','all_converged','last_frame_rel_l2_vs_truth','matrix_passes','frames_handed_their_first_iteration','sweeps_repeated')})"
