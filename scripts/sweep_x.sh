#!/bin/bash
# x-transform launch shapes, Re550 fp64, one GPU
cd "$(dirname "$0")/.."
for cfg in "128 8" "96 8" "64 8" "160 8"; do
  set -- $cfg
  echo -n "x_threads=$1 x_tk=$2 : "
  CHAN_TUNE_X_THREADS=$1 CHAN_TUNE_X_TK=$2 python bench.py --profile --steps 5 --warmup 2 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 15.0
    print('ms/step %.3f xinv %.3f z %.3f xfwd %.3f' % (d['ms_per_step'], s[5]/n, s[6]/n, s[7]/n))
except Exception as e: print('failed', e)"
done
