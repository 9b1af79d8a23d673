"""The drop-in boundary as a program: the reference's own driver (RK3.c, RK3_kernels.cu, nonLinear.c, meanUevol.c,
statistics.cu, fft.c, setUp.c — compiled where they lie by `SHIM=1 oracle/build_ref.sh`) linked against
host/channel_shim.cu, which defines the operator functions of channel.h (channel.h:175-265) on top of libchannel_b200,
in place of the reference's nine operator files.  Its dumps after two RK3 steps must equal the pure reference build's
(tests/golden) to the tolerances of the fused path: the seam carries the reference's semantics call by call."""
import os
import shutil
import subprocess
import tempfile

import numpy as np
import pytest

from tests.util import Golden, rel

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("case,tol", [("64x21x9_f64", 1e-11), ("64x21x9_f32", 2e-5)])
def test_reference_driver_on_the_shim_matches_the_reference(case, tol):
    g = Golden(case)
    binary = os.path.join(ROOT, "oracle", "_ref", "refshim_%dx%dx%d_f%d" % (g.nx, g.ny, g.nz, g.prec))
    if not os.path.exists(binary):
        pytest.skip("hybrid build %s absent (SHIM=1 oracle/build_ref.sh needs /root/reference; __graft_entry__.build() makes it)" % os.path.basename(binary))
    ddv, gg, um = g.inputs()
    d = tempfile.mkdtemp(prefix="cs", dir="/tmp")
    try:
        os.makedirs(d + "/in")
        ddv.tofile(d + "/in/ddv.bin"); gg.tofile(d + "/in/g.bin"); um.tofile(d + "/in/umean.bin")
        r = subprocess.run([binary, "steps", d + "/in", d + "/out", "2", repr(g.dt)], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-1500:])
        shape = (g.nx, g.nz, g.ny)
        for name in ("ddv", "g", "v", "dv", "R_ddv", "R_g"):
            got = np.fromfile(os.path.join(d, "out", name + ".bin"), g.cplx).reshape(shape)
            assert rel(got, g.full("steps2/" + name)) < tol, name
        um2 = np.fromfile(d + "/out/umean.bin", g.cplx)
        assert rel(um2.real, g.raw("steps2/umean").real) < (1e-12 if g.prec == 64 else 1e-6)
    finally:
        shutil.rmtree(d, ignore_errors=True)
