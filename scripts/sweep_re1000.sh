#!/bin/bash
# plane-kernel launch shapes at 1024x385x768 (x n=1024, z n=768), one GPU
cd "$(dirname "$0")/.."
for cfg in "128 8 128" "128 4 128" "128 2 128" "256 4 128" "128 4 256" "128 4 192" "64 4 256"; do
  set -- $cfg
  echo -n "x_threads=$1 x_tk=$2 z_threads=$3 : "
  CHAN_TUNE_X_THREADS=$1 CHAN_TUNE_X_TK=$2 CHAN_TUNE_Z_THREADS=$3 timeout 200 python bench.py --profile --config re1000 --steps 2 --warmup 3 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 6.0
    print('ms/step %.2f pre %.2f xinv %.2f z %.2f xfwd %.2f post %.2f' % (d['ms_per_step'], s[0]/n, s[5]/n, s[6]/n, s[7]/n, s[3]/n))
except Exception as e: print('failed', e)"
done
