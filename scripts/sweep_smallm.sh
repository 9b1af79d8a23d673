#!/bin/bash
# y-stage kernels at the per-rank pencil workload of an 8-GPU run (48x193x193 has Kx=33 of 257 kx rows)
cd "$(dirname "$0")/.."
for grid in 48,193,193 96,193,193 192,193,193; do for seg in "1 1" "1 2"; do set -- $seg; export CHAN_TUNE_POST_DEEP=$2; seg=$1;
  echo -n "grid=$grid pencil_seg=$seg deep=$CHAN_TUNE_POST_DEEP : "
  CHAN_TUNE_PENCIL_SEG=$seg python bench.py --profile --steps 5 --warmup 3 --grid $grid 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 15.0
    print('ms/step %.3f pre %.3f post %.3f scalar %.3f planes %.3f' % (d['ms_per_step'], s[0]/n, s[3]/n, s[2]/n, s[1]/n))
except Exception as e: print('failed', e)"
done; done
