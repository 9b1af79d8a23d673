#!/bin/bash
# sweep of the segment-parallel y-stage launch shapes (S*100 + variant), Re550 fp64, one GPU
cd "$(dirname "$0")/.."
for shape in 0 808 809 804 408 409 416; do
  for seg in 3; do
    echo -n "shape=$shape seg=$seg "
    CHAN_TUNE_SEG_SHAPE=$shape CHAN_TUNE_PENCIL_SEG=$seg python bench.py --profile --steps 5 --warmup 2 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); s = d['stage_ms']; n = 15.0
print('ms/step %.3f pre %.3f post %.3f (ms per substep)' % (d['ms_per_step'], s[0]/n, s[3]/n))"
  done
done
