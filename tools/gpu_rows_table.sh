#!/bin/bash
# iterations/s by row length on 8 GB matrices: two-kernel path vs fused (whole-solve kernel), pipeline depth 1 vs 3
run() {  # V sweep [env...]
  local V=$1 sweep=$2; shift 2
  local P=$((2000000000 / V))
  env "$@" timeout 200 python bench.py --npixel $P --nvoxel $V --steps 2 --warmup 3 --no-cpu-baseline --sweep $sweep 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('V=$V sweep=$sweep $*', 'it/s', round(d['value'],1), 'between_sweeps_us', round(d.get('between_sweeps_us',0),1))"
}
for V in 1024 2048 4096 8192 16384 32768; do
  run $V split A=1
  run $V fused A=1
done
for V in 4096 8192 16384; do
  run $V fused SART_FUSED_DEPTH=3
  run $V fused SART_FUSED_DEPTH=1
done
