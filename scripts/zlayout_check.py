"""The shared-memory row layouts of the z stage change no arithmetic: fields after one RK3 step must be bit-identical
with the packed layouts on and off, for several transform lengths (n = 2NZ-2)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import channel_gpu_b200 as cg
from bench import synthetic_fields

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
bad = 0
for nx, ny, nz in ((48, 17, 385), (64, 17, 193), (32, 17, 65), (96, 17, 97), (48, 17, 49), (1024, 17, 9)):
    a = run(nx, ny, nz, {"CHAN_TUNE_Z_INPACK": "0", "CHAN_TUNE_Z_FWDPACK": "0"})
    b = run(nx, ny, nz, {})
    same = all(np.array_equal(a[k], b[k]) for k in names)
    fin = all(np.isfinite(b[k]).all() for k in names)
    print("grid %dx%dx%d (z n=%d): identical=%s finite=%s" % (nx, ny, nz, 2 * nz - 2, same, fin))
    bad += (not same) or (not fin)
sys.exit(1 if bad else 0)
