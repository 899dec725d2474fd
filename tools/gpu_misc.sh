#!/bin/bash
mkdir -p gpurun_out
echo "== new tests =="
timeout 600 python -m pytest tests -m gpu -x -q -k "file_loader or reference_gpu or no_alloc" 2>&1 | tail -4
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
echo "== AUTO threshold: split vs fused by row length (8 GB matrices) =="
for V in 4096 8192 12288 16384 32768; do
  P=$((2000000000 / V))
  for sw in split fused; do
    timeout 200 python bench.py --npixel $P --nvoxel $V --steps 3 --warmup 3 --no-cpu-baseline --sweep $sw 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('V=$V P=$P', d['config']['sweep'], 'it/s', round(d['value'],1))"
  done
done
echo "== ncu full: split kernels =="
CMD="python bench.py --steps 1 --warmup 3 --iterations 2 --no-cpu-baseline --sweep split"
$CMD > gpurun_out/plain_ncu_split.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fp_rows|bp_cols" -s 6 -c 2 -f -o gpurun_out/prof_split $CMD > gpurun_out/ncu_full_split.log 2>&1
tail -3 gpurun_out/ncu_full_split.log | cut -c1-200
