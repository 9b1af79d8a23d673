"""Round-off of the register (segment-parallel) y-stage kernels against the thread-per-chain ones, one RK3 step."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import channel_gpu_b200 as cg
from bench import synthetic_fields
from tests.util import rel

names = ("ddv", "g", "v", "dv", "R_ddv", "R_g")
def run(nx, ny, nz, env):
    for k, v in env.items(): os.environ[k] = v
    ddv, gg, um = synthetic_fields(nx, ny, nz, np.complex128)
    ch = cg.Channel(nx, ny, nz, precision=8)
    ch.RKstep(ddv, gg, 1, fixed_dt=1e-3, umean=um)
    out = {k: ch.get_field(k) for k in names}
    ch.close()
    for k in env: del os.environ[k]
    return out
for ny in (97, 193, 257, 385, 513, 633):
    chain = run(24, ny, 13, {"CHAN_TUNE_PENCIL_SEG": "0", "CHAN_TUNE_POST_DEEP": "0"})
    for label, env in (("auto", {}), ("seg pre", {"CHAN_TUNE_PENCIL_SEG": "1", "CHAN_TUNE_POST_DEEP": "0"}), ("seg pre+post", {"CHAN_TUNE_PENCIL_SEG": "3"})):
        got = run(24, ny, 13, env)
        print(ny, label, " ".join("%s %.1e" % (n, rel(got[n], chain[n])) for n in names))
