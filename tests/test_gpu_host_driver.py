"""The run.conf-driven C executable (host/channel_b200) against the reference's own run of the same
case: same run.conf syntax in, field / mean-profile / statistics files out, compared with the files
the reference build wrote for 3 steps of the 64x21x9 golden case (tests/golden)."""
import os
import subprocess

import numpy as np
import pytest

from tests.util import Golden, rel

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD_DAT = os.path.join(ROOT, "tests", "golden", "64x21x9_f32_steps3_dat")


def _floats(path):
    return [[float(t) for t in line.split()] for line in open(path).read().split("\n") if line.strip()]


def test_c_driver_reproduces_reference_files(tmp_path):
    import importlib.util
    spec = importlib.util.spec_from_file_location("host_build", os.path.join(ROOT, "host", "build.py"))
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    _, exe = m.build_host()
    g = Golden("64x21x9_f32")
    ddv, gg, um = g.inputs()
    out = str(tmp_path) + "/"
    ddv.tofile(out + "ddv_in.bin"); gg.tofile(out + "g_in.bin"); um.tofile(out + "umean_in.bin")
    conf = f'''application:
{{
	NX = {g.nx};
	NY = {g.ny};
	NZ = {g.nz};
	input: {{ G = "{out}g_in.bin"; DDV = "{out}ddv_in.bin"; UMEAN = "{out}umean_in.bin"; }};
	output: {{ G = "{out}g_out.bin"; DDV = "{out}ddv_out.bin"; UMEAN = "{out}umean_out.bin"; }};
	path = "{out}";
	NSTEPS = 3; FIXED_DT = {g.dt}; PRECISION = 4;
}};
'''
    open(out + "run.conf", "w").write(conf)
    r = subprocess.run([exe, out + "run.conf"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "MEAN_PROFILE_STATISTICS" in r.stdout and "(time,counter)=" in r.stdout
    # fields and mean profile after 3 steps vs the reference build
    shape = (g.nx, g.nz, g.ny)
    assert rel(np.fromfile(out + "ddv_out.bin", np.complex64).reshape(shape), g.full("steps3/ddv")) < 5e-5
    assert rel(np.fromfile(out + "g_out.bin", np.complex64).reshape(shape), g.full("steps3/g")) < 5e-5
    assert rel(np.fromfile(out + "umean_out.bin", np.complex64).real, g.raw("steps3/umean").real) < 1e-6
    # every text file the reference wrote: same number of lines and values, values equal to print precision
    for name in sorted(os.listdir(GOLD_DAT)):
        want = _floats(os.path.join(GOLD_DAT, name)); got = _floats(out + name)
        assert len(got) == len(want), name
        for a, b in zip(got, want):
            assert len(a) == len(b), name
            a, b = np.array(a), np.array(b)
            assert np.array_equal(np.isnan(a), np.isnan(b)), name      # the -nan entries of RSTRSS.dat (Q14)
            ok = ~np.isnan(b)
            assert np.allclose(a[ok], b[ok], rtol=2e-4, atol=2e-5), (name, np.abs(a[ok] - b[ok]).max())
    # raw text shape of one file: leading blank per value, trailing blank
    line = open(out + "MEANPROFILE.dat").read().split("\n")[0]
    assert line.startswith(" ") and line.endswith(" ") and len(line.split()) == g.ny
