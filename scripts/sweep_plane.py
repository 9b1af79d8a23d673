"""Sweep the launch-shape knobs of the plane kernels (CHAN_TUNE_* env) on the bench workload:
prints the device time of the plane stage per substep for each combination."""
import itertools, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
combos = []
for zt, xp, pad, mr in itertools.product((128, 256), (0, 1, 2), (0, 1), (8, 16)):
    combos.append({"CHAN_TUNE_Z_THREADS": zt, "CHAN_TUNE_Z_XP": xp, "CHAN_TUNE_Z_PAD": pad, "CHAN_TUNE_MAX_RADIX": mr})
extra = [{"CHAN_TUNE_X_THREADS": 128}, {"CHAN_TUNE_X_TK": 4}, {"CHAN_TUNE_X_THREADS": 128, "CHAN_TUNE_X_TK": 4}, {"CHAN_TUNE_MAX_RADIX": 4}, {"CHAN_TUNE_MAX_RADIX": 6}]
res = []
for c in combos + extra:
    env = dict(os.environ); env.update({k: str(v) for k, v in c.items()})
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--profile", "--steps", "4", "--warmup", "3"] + sys.argv[1:], env=env, capture_output=True, text=True)
    try:
        j = json.loads(r.stdout.strip().splitlines()[-1])
        ms = j["stage_ms"]; n = 12
        print(json.dumps(c), "planes %.3f ms/substep  step %.2f ms" % (ms[1] / n, j["ms_per_step"]), flush=True)
    except Exception as e:
        print(json.dumps(c), "FAILED", r.stderr[-300:], flush=True)
