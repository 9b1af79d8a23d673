#!/usr/bin/env bash
# ncu evidence for one round (run under gpurun): launch list of one bench step + full
# captures of the hot kernels.  Usage: scripts/profile.sh <tag>  -> gpurun_out/<tag>_*
set -uo pipefail
TAG=${1:-r01}
CMD="python bench.py --profile --steps 1 --warmup 3"
mkdir -p gpurun_out
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
cat gpurun_out/${TAG}_plain.log | tail -1
# launch list: skip init + 3 warm-up steps, keep one timed step
ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-67} -c ${COUNT:-19} --csv \
    --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1 || echo "ncu launch list failed"
# full capture of the hot kernels (one launch each, late in the run)
ncu --set full --clock-control none --import-source on -k regex:"pencil_post|pencil_pre|zstage_kernel|xfft_inv|xfft_fwd" \
    -s ${SKIPK:-60} -c ${COUNTK:-5} -o gpurun_out/${TAG}_prof -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1 || echo "ncu full failed"
ls -la gpurun_out | grep ${TAG}
