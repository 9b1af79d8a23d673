"""Pins oracle/kmm.py (the CPU restatement) against dumps of the reference's own
CUDA build (oracle/_ref, patched per oracle/build_ref.sh) taken on a B200:
tests/golden/*.npz, made by run_golden.sh + pack_golden.py.

Tolerances.  fp32 state ("as shipped"): the y-solve operators are bit-identical
in practice; FFT-based ones differ by fp32 FFT round-off.  fp64 state ("ref64"):
1e-12, except everything downstream of setBoundaryCond, whose fp32 `expf`
(bilplacSolver_double.cu:100-217) is evaluated by CUDA's libdevice there and by
numpy here (<= 2 ulp_fp32 apart, amplified by the 2x2 wall solve).
"""
import numpy as np
import pytest

from oracle import kmm
from tests.util import Golden, rel

CASES = {"64x21x9_f32": dict(op=2e-6, bc=2e-6, step=3e-5), "64x21x9_f64": dict(op=1e-12, bc=1e-6, step=2e-6)}


@pytest.fixture(scope="module", params=sorted(CASES))
def gc(request):
    g = Golden(request.param)
    return g, kmm.Grid(g.nx, g.ny, g.nz, g.real), CASES[request.param]


def test_dealias_and_symmetry(gc):
    g, G, tol = gc
    ddv, gg, _ = g.inputs()
    assert rel(kmm.dealias(G, ddv), g.full("ops/dealias")) == 0.0
    assert rel(kmm.impose_symmetry(G, kmm.dealias(G, ddv)), g.full("ops/sym_ddv")) < tol["op"]
    assert rel(kmm.impose_symmetry(G, kmm.dealias(G, gg)), g.full("ops/sym_g")) < tol["op"]


def test_y_operators(gc):
    g, G, tol = gc
    P = g.full("ops/sym_ddv")
    b0 = kmm.BETHA[0]
    assert rel(kmm.deriv_y(G, P), g.full("ops/derivY")) < tol["op"]
    assert rel(kmm.deriv_yy(G, P), g.full("ops/derivYY")) < tol["op"]
    assert rel(kmm.helmholtz(G, P), g.full("ops/helm")) < tol["op"]
    assert rel(kmm.implicit(G, P, b0, g.dt), g.full("ops/implicit")) < tol["op"]
    v0, dv0 = kmm.calc_vdv(G, P)
    assert rel(v0, g.full("ops/v0")) < tol["op"]
    assert rel(dv0, g.full("ops/dv0")) < tol["op"]


def test_bilaplacian(gc):
    g, G, tol = gc
    a, b, c = kmm.bilaplacian(G, g.full("ops/sym_ddv"), kmm.BETHA[0], g.dt)
    assert rel(a, g.full("ops/bil_ddv")) < tol["bc"]
    assert rel(b, g.full("ops/bil_v")) < tol["bc"]
    assert rel(c, g.full("ops/bil_dv")) < tol["bc"]
    # wall conditions the correction exists for: dv(+-1) = 0, v(+-1) = 0.  They hold
    # only to ~1e-4 in the reference itself (fp32 expf and the l1-for-l2 slip, Q5).
    scale = np.abs(c).max()
    assert np.abs(c[:, :, [0, -1]]).max() < 1e-3 * scale
    assert np.abs(b[:, :, [0, -1]]).max() < 1e-3 * np.abs(b).max()


def test_pointwise_and_fft(gc):
    g, G, tol = gc
    P = g.full("ops/sym_ddv"); Gg = g.full("ops/sym_g")
    u, w = kmm.calc_uw(G, g.full("ops/dv0"), Gg)
    assert rel(u, g.full("ops/uw_u")) < tol["op"]
    assert rel(w, g.full("ops/uw_w")) < tol["op"]
    ph = kmm.to_physical(G, P)
    assert rel(ph, g.raw("ops/phys")) < tol["op"]
    assert rel(kmm.to_spectral(G, ph), g.full("ops/roundtrip")) < tol["op"]
    H = kmm.convolution(G, g.full("ops/uw_u"), g.full("ops/v0"), g.full("ops/uw_w"))
    for h, n in zip(H, ("conv_wx", "conv_wy", "conv_wz")):
        assert rel(h, g.full("ops/" + n)) < tol["op"]
    hv, hg = kmm.calc_hvg(G, g.full("ops/conv_wx"), g.full("ops/conv_wy"), g.full("ops/conv_wz"))
    assert rel(hv, g.full("ops/hvg_hv")) < tol["op"]
    assert rel(hg, g.full("ops/hvg_hg")) < tol["op"]


def test_rk_pointwise(gc):
    g, G, tol = gc
    P = g.full("ops/sym_ddv"); Gg = g.full("ops/sym_g")
    pw = kmm.rk_step_1(G, P, kmm.deriv_yy(G, P), Gg, g.dt, 1)
    gw = kmm.rk_step_1(G, Gg, kmm.deriv_yy(G, Gg), P, g.dt, 1)
    assert rel(pw, g.full("ops/rk1_ddv_w")) < tol["op"]
    assert rel(gw, g.full("ops/rk1_g_w")) < tol["op"]
    assert rel(kmm.rk_step_2(G, g.full("ops/rk1_ddv_w"), Gg, g.dt, 1), g.full("ops/rk2_ddv_w")) < tol["op"]


def _check_steps(g, G, sub, n, fixed, tol):
    ddv, gg, um = g.inputs()
    S = kmm.State(G, ddv, gg, um).run(n, fixed)
    for name, arr in (("ddv", S.ddv), ("g", S.g), ("v", S.v), ("dv", S.dv), ("R_ddv", S.R_ddv), ("R_g", S.R_g)):
        assert rel(arr, g.full(f"{sub}/{name}")) < tol, name
    assert rel(S.mean.u, g.raw(f"{sub}/umean").real) < 1e-6
    ref_dt = g.raw(f"{sub}/dt")[0]                    # calcDt's own value at the last step
    assert abs(float(S.last_dt_c) - ref_dt) <= 2e-6 * ref_dt
    return S


def test_rk3_steps(gc):
    g, G, tol = gc
    for sub in g.meta["subs"]:
        if sub.startswith("steps") and sub[5:].isdigit():
            _check_steps(g, G, sub, int(sub[5:]), g.dt, tol["step"])


def test_rk3_free_running_dt():
    g = Golden("64x21x9_f32")
    G = kmm.Grid(g.nx, g.ny, g.nz, g.real)
    _check_steps(g, G, "steps3_freedt", 3, None, 3e-5)


def test_rk3_radix3_grid():
    """NX = 3*64, Nz = 3*8, NX != NY, NZ*NY % 16 != 0 (patches P1, P8)."""
    g = Golden("192x19x13_f32")
    G = kmm.Grid(g.nx, g.ny, g.nz, g.real)
    _check_steps(g, G, "steps2", 2, g.dt, 1e-4)
