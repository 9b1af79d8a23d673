#!/bin/bash
# exchange variants at N GPUs, Re_tau=1000 (run under gpurun --gpus N)
cd "$(dirname "$0")/.."
N=${N:-8}; TAG=${TAG:-r02_n8x}; OUT=gpurun_out/$TAG; mkdir -p $OUT
# VARIANTS: ';'-separated lists of VAR=value (default: the four exchange designs)
IFS=';' read -ra VARS <<< "${VARIANTS:-CHAN_TUNE_XCHG_PIPE=0;CHAN_TUNE_XCHG_PIPE=1;CHAN_TUNE_PRE_DIRECT=1 CHAN_TUNE_FWD_DIRECT=1;CHAN_TUNE_XCHG_PIPE=0 CHAN_TUNE_FWD_DIRECT=1}"
for V in "${VARS[@]}"; do
  name=$(echo $V | tr ' =' '__')
  env $V timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $N --config ${CFG:-re1000} --steps 5 --warmup 3 > $OUT/$name.json 2> $OUT/$name.err
  tail -1 $OUT/$name.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
s=d['stages']
print(d['config']['knobs'], 'ms/step %.2f e2e %.2f parity %.2e nvlink %.0f GB/s' % (d['ms_per_step'], d['e2e']['value'], d['parity_vs_golden']['worst_error_over_tolerance'], d['nvlink']['achieved_GBps']))
print('   ', {k:round(v['ms_per_substep'],3) for k,v in s.items() if isinstance(v,dict)}, round(s['residual_ms_per_step'],3))
" || tail -3 $OUT/$name.err
done
