#!/bin/bash
# First GPU session: box facts, GPU tests, bench, launch list, kernel-variant sweep, read-bandwidth calibration.
mkdir -p gpurun_out
{
  echo "== box =="; nvidia-smi --query-gpu=name,memory.total,memory.used,clocks.max.sm,clocks.max.mem,power.limit --format=csv
  nproc; free -g | head -2; lscpu | grep -E "Model name|Socket|Core|Thread" 
} > gpurun_out/box.txt 2>&1
echo "== pytest -m gpu ==" 
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 | tee gpurun_out/pytest_gpu.txt
echo "== smoke =="
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5 | tee gpurun_out/smoke.txt
echo "== membw =="
timeout 120 tools/bin/membw 16 2>&1 | tee gpurun_out/membw.txt
echo "== bench =="
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
echo "== variants =="
for fv in 0 1 2 3; do
  SART_FP_VARIANT=$fv timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); k = d['roofline']['kernels']
        print('FP variant $fv', 'value', round(d['value'],2), 'fp GB/s', round(k['fp_rows_kernel']['gbs'],1), 'bp GB/s', round(k['bp_cols_kernel']['gbs'],1))
" | tee -a gpurun_out/variants.txt
done
for bv in 1 2 3; do
  SART_BP_VARIANT=$bv timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); k = d['roofline']['kernels']
        print('BP variant $bv', 'value', round(d['value'],2), 'fp GB/s', round(k['fp_rows_kernel']['gbs'],1), 'bp GB/s', round(k['bp_cols_kernel']['gbs'],1))
" | tee -a gpurun_out/variants.txt
done
for s in 6 24 48; do
  SART_BP_SLABS=$s timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); k = d['roofline']['kernels']
        print('BP slabs $s', 'value', round(d['value'],2), 'fp GB/s', round(k['fp_rows_kernel']['gbs'],1), 'bp GB/s', round(k['bp_cols_kernel']['gbs'],1))
" | tee -a gpurun_out/variants.txt
done
echo "== ncu launch list =="
python bench.py --steps 1 --warmup 3 --iterations 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 3 --iterations 3 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
echo done
