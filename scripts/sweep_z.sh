#!/bin/bash
# z-stage launch shapes, Re550 fp64, one GPU: threads per block x x-pairs per block
cd "$(dirname "$0")/.."
for cfg in "0 0" "256 2" "192 2" "96 1" "160 1"; do
  set -- $cfg
  echo -n "z_threads=$1 z_xp=$2 : "
  CHAN_TUNE_Z_THREADS=$1 CHAN_TUNE_Z_XP=$2 python bench.py --profile --steps 5 --warmup 2 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 15.0
    print('ms/step %.3f xinv %.3f z %.3f xfwd %.3f' % (d['ms_per_step'], s[5]/n, s[6]/n, s[7]/n))
except Exception as e: print('failed', e)"
done
