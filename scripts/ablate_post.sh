#!/bin/bash
# cumulative time of the chain post kernel's passes (kernel returns after pass k; results are garbage)
# needs an ablation build first:  CHAN_ABLATE=1 python -m channel_gpu_b200.build --force   (and a normal rebuild afterwards)
cd "$(dirname "$0")/.."
for k in 1 2 3 4 5 6 0; do
  echo -n "post_stop=$k : "
  CHAN_TUNE_POST_STOP=$k python bench.py --allow-knobs --profile --steps 4 --warmup 2 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 12.0
    print('post %.3f ms/substep' % (s[3]/n))
except Exception as e: print('failed', e)"
done
