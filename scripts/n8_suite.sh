#!/bin/bash
# N-GPU measurements of one round (run under `gpurun --gpus N`, N=8 by default): parity of all ranks on the golden cases,
# Re_tau=1000, Re_tau=2000 (BASELINE.json configs[3]), Re_tau=550 with the library defaults.  One JSON line per run in gpurun_out/$TAG/.
cd "$(dirname "$0")/.."
TAG=${TAG:-r02_n8}; N=${N:-8}; OUT=gpurun_out/$TAG; mkdir -p $OUT
run() {  # name, env..., -- bench args
  local name=$1; shift
  local envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 \
      bench.py --gpus $N "$@" > $OUT/$name.json 2> $OUT/$name.err || { echo "$name FAILED"; tail -5 $OUT/$name.err; }
  python - $OUT/$name.json $name <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().split("\n")[-1])
    s = d["stages"]
    print("%-28s %8.3f ms/step  e2e %.2f steps/s  parity %s  pre %.3f planes %.3f (xi %.3f z %.3f xf %.3f) post %.3f scalar %.3f resid %.3f  nvlink %.0f GB/s" % (
        sys.argv[2], d["ms_per_step"], d["e2e"]["value"], d.get("parity_vs_golden", {}).get("worst_error_over_tolerance"),
        s["pencil_pre"]["ms_per_substep"], s["plane_stage_total"]["ms_per_substep"], s["xfft_inv"]["ms_per_substep"], s["zstage"]["ms_per_substep"],
        s["xfft_fwd"]["ms_per_substep"], s["pencil_post"]["ms_per_substep"], s["scalar"]["ms_per_substep"], s["residual_ms_per_step"], d.get("nvlink", {}).get("achieved_GBps") or 0))
except Exception as e:
    print(sys.argv[2], "no result:", e)
PY
}
[ "${CHECK:-1}" = 1 ] && timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29654 tests/mgpu_check.py 2>&1 | grep "parity\|rank" | sort | head -12 | tee $OUT/mgpu_check.log
[ "${SKIP1000:-0}" = 1 ] || run re1000_default -- --config re1000 --steps ${STEPS:-10} --warmup 3
if [ "${FULL:-1}" != 0 ]; then
run re2000_default -- --config re2000 --steps 3 --warmup 3
fi
if [ "${FULL:-1}" = 1 ]; then
run re550_default -- --config re550 --steps 20 --warmup 3
fi
nvidia-smi --query-gpu=index,memory.used,memory.total --format=csv | tee $OUT/mem.txt
