"""Checkpoint / restart: the reference restarts from its end-of-run field files and loses
R_phi, R_g, dt and the step count (main.c:139, RK3.c:114); here a run continued from a checkpoint must be the
uninterrupted run bit for bit — through the C ABI (chan_checkpoint_save / _load) and through the run.conf-driven
C program (application.checkpoint.{EVERY,FILE,RESTART})."""
import os
import subprocess

import numpy as np
import pytest

from tests.util import Golden

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ("ddv", "g", "v", "dv", "R_ddv", "R_g")


@pytest.mark.parametrize("case", ["64x21x9_f64", "64x21x9_f32"])
def test_two_plus_two_steps_equal_four(case):
    import channel_gpu_b200 as cg
    g = Golden(case)
    ddv, gg, um = g.inputs()
    prec = 4 if g.prec == 32 else 8
    a = cg.Channel(g.nx, g.ny, g.nz, precision=prec)
    a.set_fields(ddv, gg); a.set_umean(um); a.rk_init()       # free-running dt: the lagged dt must survive the restart
    a.rk_step(2); a.sync()
    blob = a.checkpoint_save()
    a.rk_step(2); a.sync()
    want = {k: a.get_field(k) for k in NAMES}; want_um = a.get_umean(); want_sc = a.scalars()
    a.close()
    b = cg.Channel(g.nx, g.ny, g.nz, precision=prec)          # a fresh handle: nothing but the blob carries over
    b.checkpoint_load(blob)
    assert b.scalars()["counter"] == 2
    b.rk_step(2); b.sync()
    for k in NAMES:
        assert np.array_equal(b.get_field(k), want[k]), k
    assert np.array_equal(b.get_umean(), want_um)
    sc = b.scalars()
    for k in ("dt", "dt_next", "dt_c", "counter", "time", "utau"):
        assert sc[k] == want_sc[k], k
    assert sc["counter"] == 4 and sc["time"] > 0
    # a blob of another grid is refused
    c = cg.Channel(48, 25, 13, precision=prec)
    with pytest.raises(cg.ChannelError):
        c.checkpoint_load(blob)
    c.close(); b.close()


def test_c_driver_restart_continues_the_run(tmp_path):
    import importlib.util
    spec = importlib.util.spec_from_file_location("host_build", os.path.join(ROOT, "host", "build.py"))
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    _, exe = m.build_host()
    g = Golden("64x21x9_f32")
    ddv, gg, um = g.inputs()
    out = str(tmp_path) + "/"
    ddv.tofile(out + "ddv_in.bin"); gg.tofile(out + "g_in.bin"); um.tofile(out + "umean_in.bin")

    def conf(tag, nsteps, extra):
        os.makedirs(out + tag, exist_ok=True)
        text = f'''application:
{{
	NX = {g.nx}; NY = {g.ny}; NZ = {g.nz};
	input: {{ G = "{out}g_in.bin"; DDV = "{out}ddv_in.bin"; UMEAN = "{out}umean_in.bin"; }};
	output: {{ G = "{out}{tag}/G.h5"; DDV = "{out}{tag}/ddV.h5"; UMEAN = "{out}{tag}/umean.bin"; }};
	path = "{out}{tag}/";
	NSTEPS = {nsteps}; PRECISION = 4;
	{extra}
}};
'''
        open(out + tag + ".conf", "w").write(text)
        r = subprocess.run([exe, out + tag + ".conf"], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        return r.stdout

    conf("whole", 5, "")
    conf("first", 3, f'checkpoint: {{ EVERY = 3; FILE = "{out}c.ckpt"; }};')
    log = conf("second", 5, f'checkpoint: {{ RESTART = "{out}c.ckpt"; }};')
    assert "Restarted from" in log and "counter)=(" in log
    for f in ("G.h5", "ddV.h5", "umean.bin"):           # HDF5 containers (host/h5min.c) and the raw mean profile
        assert open(out + "whole/" + f, "rb").read() == open(out + "second/" + f, "rb").read(), f
    # the per-step text files of the continued run are the tail of the uninterrupted run's
    whole = open(out + "whole/MEANPROFILE.dat").read().split("\n")
    tail = open(out + "second/MEANPROFILE.dat").read().split("\n")
    assert len(whole) == 6 and len(tail) == 3 and whole[3:] == tail
