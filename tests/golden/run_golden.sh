#!/usr/bin/env bash
# Runs on the GPU box (gpurun): feeds the seeded inputs of make_inputs.py to the
# reference builds under oracle/_ref/ and leaves the raw dumps in
# gpurun_out/golden/<case>/.  tests/golden/pack_golden.py turns them into the
# committed .npz fixtures.  Nothing here reads /root/reference.
set -uo pipefail
cd "$(dirname "$0")/../.."
OUT=gpurun_out/golden
mkdir -p "$OUT"
DT=0.005
run_case() {  # NX NY NZ PREC "ops" "steps list"
  local nx=$1 ny=$2 nz=$3 prec=$4 doops=$5 steps=$6
  local bin=oracle/_ref/ref_${nx}x${ny}x${nz}_f${prec}
  local c=$OUT/${nx}x${ny}x${nz}_f${prec}
  [ -x "$bin" ] || { echo "missing $bin"; return 1; }
  python tests/golden/make_inputs.py $nx $ny $nz $c/in $prec
  if [ "$doops" = 1 ]; then
    timeout 300 $bin ops $c/in $c/ops $DT > $c/ops.log 2>&1 || { echo "ops FAILED $c"; tail -5 $c/ops.log; }
  fi
  for n in $steps; do
    timeout 300 $bin steps $c/in $c/steps$n $n $DT > $c/steps$n.log 2>&1 || { echo "steps$n FAILED $c"; tail -5 $c/steps$n.log; }
  done
  rm -rf $c/in
  echo "case $c done"
}
run_case 64 21 9 32 1 "1 2 3"
run_case 64 21 9 64 1 "2"
run_case 192 19 13 32 0 "2"
# free-running dt (calcDt, RK3.c:64-109) on the small case, for the dt comparison
c=$OUT/64x21x9_f32
python tests/golden/make_inputs.py 64 21 9 $c/in 32
timeout 300 oracle/_ref/ref_64x21x9_f32 steps $c/in $c/steps3_freedt 3 0 > $c/steps3_freedt.log 2>&1 || echo "free dt FAILED"
rm -rf $c/in
# reference timing at the BASELINE grids that were built
for b in oracle/_ref/ref_128x97x65_f32 oracle/_ref/ref_128x97x65_f64 oracle/_ref/ref_384x193x193_f32 oracle/_ref/ref_384x193x193_f64; do
  [ -x $b ] && { timeout 600 $b time 10 3 2>&1 | grep REF_TIME || echo "time FAILED $b"; } | tee -a $OUT/ref_time.log
done
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv | tee $OUT/gpu.txt
du -sh $OUT
