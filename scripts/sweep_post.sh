#!/bin/bash
# chain post kernel: L2 prefetch on/off x block size, Re550 fp64, one GPU
cd "$(dirname "$0")/.."
for cfg in "0 128" "1 128" "2 128" "3 128" "2 64"; do
  set -- $cfg
  echo -n "post_pf=$1 post_threads=$2 : "
  CHAN_TUNE_POST_PF=$1 CHAN_TUNE_POST_THREADS=$2 python bench.py --profile --steps 5 --warmup 2 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 15.0
    print('ms/step %.3f pre %.3f z %.3f post %.3f' % (d['ms_per_step'], s[0]/n, s[6]/n, s[3]/n))
except Exception as e: print('failed', e)"
done
