"""CPU restatement (numpy) of the hot path of Nasrollah/CHANNEL_GPU: one
Kim-Moin-Moser (phi = lap v, g = omega_y) RK3 step.

TEST INFRASTRUCTURE ONLY.  Imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py, never by the product
(channel_gpu_b200/ fails loudly without its CUDA library).

Every function names the reference lines it restates (paths relative to
/root/reference/src).  Arrays use the reference layout [i][k][j] = [kx index]
[kz][y], y fastest (RK3_kernels.cu:16-18), full NX x NZ x NY (no compaction), so
they compare 1:1 with the raw dumps of the reference build (oracle/_ref, see
tests/golden/).

Precision: `real` is the state type: np.float32 mirrors the reference as
shipped (fp32 at rest, fp64 inside the y-solves); np.float64 mirrors "ref64",
the same text with the state type widened by sed (oracle/build_ref.sh).  In both,
literals keep their f suffix, i.e. RK coefficients, pi and DELTA_Y are
float32-VALUED constants.

Parity status: pinned against dumps of the patched reference build run on a
B200 (tests/golden/*.npz, tests/test_oracle_golden.py).  The reference ships no
tests or golden vectors of its own.

Known, documented deviations from the reference text:
  Q7  max|u| for the CFL estimate is taken over valid physical points only (the
      reference's isamax also scans the two pad floats per row, fft.c:147-158);
  Q13 implicitStepMeanU's reads of u[-1], u[NY] (meanUevol.c:415) are not made.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32
F64 = np.float64

# RK3.c:14, RK3_kernels.cu:24-25,113, meanUevol.c:35-38 (float literals)
ALPHA = [F32(29.0) / F32(96.0), F32(-3.0) / F32(40.0), F32(1.0) / F32(6.0)]
BETHA = [F32(37.0) / F32(160.0), F32(5.0) / F32(24.0), F32(1.0) / F32(6.0)]
GAMMAD = [F32(8.0) / F32(15.0), F32(5.0) / F32(12.0), F32(3.0) / F32(4.0)]
DSEDA = [F32(0.0), F32(-17.0) / F32(60.0), F32(-5.0) / F32(12.0)]


class Grid:
    """channel.h:42-62 constants for one grid and state type."""

    def __init__(self, nx, ny, nz, real=np.float64, reynolds=3250.0, qvelocity=1.8, i_range=None):
        self.nx, self.ny, self.nz = int(nx), int(ny), int(nz)
        self.real = np.dtype(real).type
        self.cplx = np.complex64 if self.real is F32 else np.complex128
        R = self.real
        self.nzp = 2 * self.nz - 2                       # physical z points (fft.c:17)
        self.N2 = self.nx * self.nzp                     # convolution_kernels.cu:93
        self.LY = R(F32(2.0))                            # channel.h:50
        self.DELTA_Y = R(F32(2.0) / F32(self.ny - 1))    # channel.h:51 (float division)
        self.PI2 = R(F32(2.0) * F32(3.14159265))         # channel.h:53
        self.LX = R(F32(1.0)) * self.PI2                 # channel.h:56
        self.LZ = R(F32(0.5)) * self.PI2                 # channel.h:57
        self.REYNOLDS = R(F32(reynolds))                 # channel.h:61
        self.QVELOCITY = R(F32(qvelocity))               # channel.h:62
        # i_range = (i0, i1): this object describes the kx slab of one rank (IGLOBAL, channel.h:76);
        # pencil operators then act on arrays [i1-i0][NZ][NY].  The FFT operators need the full grid.
        i = np.arange(self.nx) if i_range is None else np.arange(i_range[0], i_range[1])
        self.kx_int = np.where(i < self.nx // 2, i, i - self.nx)          # RK3_kernels.cu:45
        self.kz_int = np.arange(self.nz)
        # wavenumbers in the state type (RK3_kernels.cu:45-54) ...
        self.k1 = (self.PI2 / self.LX) * self.kx_int.astype(R)
        self.k3 = (self.PI2 / self.LZ) * self.kz_int.astype(R)
        self.kk = (self.k1[:, None] * self.k1[:, None] + self.k3[None, :] * self.k3[None, :]).astype(R)
        # ... and in double, starting from the same real-typed quotient (hemholzt_nu_double.cu:121-131)
        self.k1d = F64(self.PI2 / self.LX) * self.kx_int.astype(F64)
        self.k3d = F64(self.PI2 / self.LZ) * self.kz_int.astype(F64)
        self.kkd = self.k1d[:, None] ** 2 + self.k3d[None, :] ** 2
        self.active = (np.abs(self.kx_int)[:, None] <= self.nx // 3) & (self.kz_int[None, :] <= self.nzp // 3)  # dealias.cu:37
        self._coef = None

    # -- mesh ---------------------------------------------------------------
    def ymesh_d(self, j):
        """Fmesh(j*DELTA_Y-1.0) as the device kernels evaluate it: product in the
        state type, then double (channel.h:21; derivatives_nu_double.cu:60)."""
        j = np.asarray(j)
        x = (j.astype(self.real) * self.DELTA_Y).astype(F64) - 1.0
        return np.tanh(2.0 * x) / np.tanh(2.0)

    def ymesh_h(self, j):
        """Fmesh(j*DELTA_Y-1.0f) as the host mean-flow code evaluates it
        (meanUevol.c:207,230): argument in the state type, tanh in double."""
        j = np.asarray(j)
        x = (j.astype(self.real) * self.DELTA_Y - self.real(1.0)).astype(F64)
        return np.tanh(2.0 * x) / np.tanh(2.0)

    def coef(self):
        """All y-dependent double coefficients used by the fp64 solvers."""
        if self._coef is not None:
            return self._coef
        ny = self.ny
        j = np.arange(ny)
        y = self.ymesh_d
        a = y(j + 1) - y(j)
        b = y(j - 1) - y(j)
        c = {}
        # first derivative, interior (derivatives_nu_double.cu:60-65, 227-232)
        c["A1"] = 2.0 * (2.0 * a * b * b - b * b * b) / ((a * a - a * b) * (a - b) * (a - b))
        c["B1"] = 2.0 * (2.0 * b * a * a - a * a * a) / ((b * b - a * b) * (a - b) * (a - b))
        c["C1"] = -c["A1"] - c["B1"]
        c["al1"] = b * b / ((a - b) * (a - b))
        c["be1"] = a * a / ((a - b) * (a - b))
        # walls (derivatives_nu_double.cu:85-113, 250-274)
        a0 = y(1) - y(0); b0 = y(2) - y(1)
        aN = y(ny - 2) - y(ny - 1); bN = y(ny - 3) - y(ny - 2)
        c["w0"] = (-(3.0 * a0 + 2.0 * b0) / (a0 * a0 + a0 * b0), ((a0 + b0) * (2.0 * b0 - a0)) / (a0 * b0 * b0), a0 * a0 / (b0 * b0 * a0 + b0 * b0 * b0), (a0 + b0) / b0)
        c["wN"] = (-(3.0 * aN + 2.0 * bN) / (aN * aN + aN * bN), ((aN + bN) * (2.0 * bN - aN)) / (aN * bN * bN), aN * aN / (bN * bN * aN + bN * bN * bN), (aN + bN) / bN)
        # second-derivative compact weights (hemholzt_nu_double.cu:140-145)
        den = a * a * a - b * b * b - 4.0 * a * a * b + 4.0 * b * b * a
        c["A2"] = -12.0 * b / den
        c["B2"] = 12.0 * a / den
        c["C2"] = -c["A2"] - c["B2"]
        den2 = a * a * a - 4.0 * a * a * b + 4.0 * a * b * b - b * b * b
        c["al2"] = -(-b * b * b - a * b * b + a * a * b) / den2
        c["be2"] = -(a * a * a + b * a * a - b * b * a) / den2
        c["y"] = y(j)
        self._coef = c
        return c


# ---------------------------------------------------------------------------
# tridiagonal solve (stands in for cusparseZgtsvStridedBatch: exact up to fp64
# round-off; derivatives_nu_double.cu:406)
# ---------------------------------------------------------------------------
def thomas(l, c, u, d):
    """Solve l[j] x[j-1] + c[j] x[j] + u[j] x[j+1] = d[j] along the last axis.
    l, c, u broadcast against d (real double), d complex128."""
    n = d.shape[-1]
    shape = d.shape
    l = np.broadcast_to(l, shape); c = np.broadcast_to(c, shape); u = np.broadcast_to(u, shape)
    cp = np.empty(shape, F64); dp = np.empty(shape, np.complex128)
    cp[..., 0] = u[..., 0] / c[..., 0]
    dp[..., 0] = d[..., 0] / c[..., 0]
    for j in range(1, n):
        m = 1.0 / (c[..., j] - l[..., j] * cp[..., j - 1])
        cp[..., j] = u[..., j] * m
        dp[..., j] = (d[..., j] - l[..., j] * dp[..., j - 1]) * m
    x = np.empty(shape, np.complex128)
    x[..., n - 1] = dp[..., n - 1]
    for j in range(n - 2, -1, -1):
        x[..., j] = dp[..., j] - cp[..., j] * x[..., j + 1]
    return x


def _shift(f, s):
    """f[..., j+s] with zeros outside (the reference reads sf[0], sf[NY+1] there
    and discards the result; derivatives_nu_double.cu:76-78)."""
    out = np.zeros_like(f)
    if s == 1:
        out[..., :-1] = f[..., 1:]
    elif s == -1:
        out[..., 1:] = f[..., :-1]
    return out


# ---------------------------------------------------------------------------
# L2 numerical primitives
# ---------------------------------------------------------------------------
def deriv_y(G: Grid, f):
    """deriv_Y_HO_double (derivatives_nu_double.cu:390-415; kernels :32-120,
    :209-285): compact 4th-order d/dy, fp64 inside, result cast to the state type."""
    c = G.coef(); ny = G.ny
    fd = f.astype(np.complex128)
    rhs = c["A1"] * _shift(fd, 1) + c["C1"] * fd + c["B1"] * _shift(fd, -1)
    A, B, C, al0 = c["w0"]
    rhs[..., 0] = A * fd[..., 0] + B * fd[..., 1] + C * fd[..., 2]
    A, B, C, alN = c["wN"]
    rhs[..., ny - 1] = A * fd[..., ny - 1] + B * fd[..., ny - 2] + C * fd[..., ny - 3]
    l = c["be1"].copy(); d = np.ones(ny); u = c["al1"].copy()
    u[0] = al0; l[0] = 0.0
    l[ny - 1] = alN; u[ny - 1] = 0.0
    return thomas(l, d, u, rhs).astype(G.cplx)


def deriv_yy(G: Grid, f):
    """deriv_YY_HO_double (derivatives_nu_double.cu:419-451): d/dy applied twice,
    with the cast to the state type in between (Q4)."""
    return deriv_y(G, deriv_y(G, f))


def _helm_rhs(G, f):
    """rhs_A_kernel (hemholzt_nu_double.cu:36-96 == implicitStep_nu_double.cu:36-96)."""
    c = G.coef()
    fd = f.astype(np.complex128)
    rhs = c["al2"] * _shift(fd, 1) + fd + c["be2"] * _shift(fd, -1)
    rhs[..., 0] = 0.0
    rhs[..., G.ny - 1] = 0.0
    return rhs


def _wall_rows(l, d, u, ny):
    """Dirichlet rows and decoupling (hemholzt_nu_double.cu:167-185)."""
    l[..., 0] = 0.0; d[..., 0] = 1.0; u[..., 0] = 0.0
    l[..., 1] = 0.0
    l[..., ny - 1] = 0.0; d[..., ny - 1] = 1.0; u[..., ny - 1] = 0.0
    u[..., ny - 2] = 0.0


def helmholtz(G: Grid, f):
    """hemholztSolver_double (hemholzt_nu_double.cu:228-251; setDiagkernel :98-195):
    (D2 - kk) v = f with v(+-1) = 0."""
    c = G.coef(); kk = G.kkd[:, :, None]
    u = c["A2"] - kk * c["al2"]
    d = c["C2"] - kk + 0.0 * c["al2"]
    l = c["B2"] - kk * c["be2"]
    _wall_rows(l, d, u, G.ny)
    return thomas(l, d, u, _helm_rhs(G, f)).astype(G.cplx)


def implicit(G: Grid, f, betha, dt):
    """implicitSolver_double (implicitStep_nu_double.cu:227-247; setDiagkernel
    :98-195): (1 - D (D2 - kk)) u = f, D = nu*betha*dt, u(+-1)=0.  `dt*betha` is
    formed in the state type (:235), nu = 1.0/REYNOLDS in double (:133)."""
    c = G.coef(); kk = G.kkd[:, :, None]
    bdt = F64(G.real(dt) * G.real(betha))
    nu = 1.0 / F64(G.REYNOLDS)
    D = nu * bdt
    l = c["be2"] - D * c["B2"] + D * kk * c["be2"]
    d = 1.0 - D * c["C2"] + D * kk * 1.0
    u = c["al2"] - D * c["A2"] + D * kk * c["al2"]
    _wall_rows(l, d, u, G.ny)
    return thomas(l, d, u, _helm_rhs(G, f)).astype(G.cplx)


def _expf(x):
    """expf(double) of bilplacSolver_double.cu:100-217: argument and result in float (Q5)."""
    with np.errstate(over="ignore", under="ignore"):
        return np.exp(np.asarray(x, F64).astype(F32)).astype(F64)


def _bc_functions(l1, l2, D, y):
    """v1,v2,dv1,dv2,ddv1,ddv2 exactly as written at bilplacSolver_double.cu:174-217
    (the l2 lines use exp(-l1*(1+y)) in their first bracket, Q5)."""
    E1m = _expf(-l1 * (1.0 - y)); E1p = np.exp(-l1 * (1.0 + y))
    E2m = _expf(-l2 * (1.0 - y)); E2p = np.exp(-l2 * (1.0 + y))
    p1 = 1.0 + np.exp(-l1 * 2.0); m1 = 1.0 - np.exp(-l1 * 2.0)
    p2 = 1.0 + np.exp(-l2 * 2.0); m2 = 1.0 - np.exp(-l2 * 2.0)
    v1 = 0.5 * D * ((E1m + E1p) / p1 - (E1m - E1p) / m1)
    v1 = v1 + 0.5 * D * (-(E2m + E1p) / p2 + (E2m - E2p) / m2)
    v2 = 0.5 * D * ((E1m + E1p) / p1 + (E1m - E1p) / m1)
    v2 = v2 + 0.5 * D * (-(E2m + E1p) / p2 - (E2m - E2p) / m2)
    dv1 = 0.5 * D * l1 * ((E1m - E1p) / p1 - (E1m + E1p) / m1)
    dv1 = dv1 + 0.5 * D * l2 * (-(E2m - E1p) / p2 + (E2m + E2p) / m2)
    dv2 = 0.5 * D * l1 * ((E1m - E1p) / p1 + (E1m + E1p) / m1)
    dv2 = dv2 + 0.5 * D * l2 * (-(E2m - E1p) / p2 - (E2m + E2p) / m2)
    ddv1 = 0.5 * ((E1m + E1p) / p1 - (E1m - E1p) / m1)
    ddv2 = 0.5 * ((E1m + E1p) / p1 + (E1m - E1p) / m1)
    return v1, v2, dv1, dv2, ddv1, ddv2


def set_boundary_cond(G: Grid, ddv, v, dv, betha, dt):
    """setBoundaryCond (bilplacSolver_double.cu:6-296): adds the two homogeneous
    solutions so that dv(+-1)=0; zeroes mode (0,0)."""
    kk = G.kkd[:, :, None]
    with np.errstate(all="ignore"):
        l1 = np.sqrt(kk + F64(G.REYNOLDS) / (F64(G.real(dt)) * F64(G.real(betha))))
        l2 = np.sqrt(kk)
        D = 1.0 / (l1 * l1 - l2 * l2)
        _, _, dv1_p, dv2_p, _, _ = _bc_functions(l1, l2, D, 1.0)
        _, _, dv1_m, dv2_m, _, _ = _bc_functions(l1, l2, D, -1.0)
        dvd = dv.astype(np.complex128)
        dv_m = dvd[:, :, 0:1]; dv_p = dvd[:, :, G.ny - 1:G.ny]
        det = dv1_m * dv2_p - dv1_p * dv2_m
        c1 = -(dv2_p * dv_m - dv2_m * dv_p) / det
        c2 = -(-dv1_p * dv_m + dv1_m * dv_p) / det
        y = G.coef()["y"][None, None, :]
        v1, v2, dv1, dv2, ddv1, ddv2 = _bc_functions(l1, l2, D, y)
        ddv_o = ddv.astype(np.complex128) + c1 * ddv1 + c2 * ddv2
        dv_o = dvd + c1 * dv1 + c2 * dv2
        v_o = v.astype(np.complex128) + c1 * v1 + c2 * v2
    if G.kx_int[0] == 0:
        ddv_o[0, 0, :] = 0.0; dv_o[0, 0, :] = 0.0; v_o[0, 0, :] = 0.0
    return ddv_o.astype(G.cplx), v_o.astype(G.cplx), dv_o.astype(G.cplx)


def bilaplacian(G: Grid, ddv, betha, dt):
    """bilaplaSolver_double (bilplacSolver_double.cu:320-348) -> (ddv, v, dv)."""
    ddv = implicit(G, ddv, betha, dt)
    v = helmholtz(G, ddv)
    dv = deriv_y(G, v)
    if G.real(dt) > 1e-8:                               # bilplacSolver_double.cu:314 (Q6)
        ddv, v, dv = set_boundary_cond(G, ddv, v, dv, betha, dt)
    return ddv, v, dv


def calc_vdv(G: Grid, ddv):
    """calcVdV (RK3.c:52-62)."""
    v = helmholtz(G, ddv)
    return v, deriv_y(G, v)


# ---------------------------------------------------------------------------
# L3 physics operators (state-type arithmetic)
# ---------------------------------------------------------------------------
def dealias(G: Grid, f):
    """dealiaskernel (dealias.cu:9-45): 2/3 truncation in place."""
    return np.where(G.active[:, :, None], f, 0).astype(G.cplx)


def _parts(f):
    return f.real.copy(), f.imag.copy()


def _mk(G, re, im):
    out = np.empty(re.shape, G.cplx)
    out.real = re; out.imag = im
    return out


def calc_uw(G: Grid, dv, g):
    """calcUWkernel (nonLinear_kernels.cu:8-92): (dv, g) -> (u, w); mode (0,0) zeroed."""
    R = G.real
    k1 = G.k1[:, None, None]; k3 = G.k3[None, :, None]
    kk = G.kk.copy()
    if G.kx_int[0] == 0:
        kk[0, 0] = R(1.0)
    kk = kk[:, :, None]
    fx, fy = _parts(dv); gx, gy = _parts(g)
    fx = -fx; fy = -fy
    u1x = -(k1 * fy + k3 * gy); u1y = k1 * fx + k3 * gx
    u1x = -u1x / kk; u1y = -u1y / kk
    u3x = -(k3 * fy - k1 * gy); u3y = k3 * fx - k1 * gx
    u3x = -u3x / kk; u3y = -u3y / kk
    u = _mk(G, u1x, u1y); w = _mk(G, u3x, u3y)
    if G.kx_int[0] == 0:
        u[0, 0, :] = 0; w[0, 0, :] = 0
    return u, w


def calc_omega(G: Grid, wx, wz, u, v, w):
    """calcOmegakernel (convolution_kernels.cu:7-66); on entry wx = dw/dy, wz = du/dy."""
    k1 = G.k1[:, None, None]; k3 = G.k3[None, :, None]
    u1x, u1y = _parts(u); u2x, u2y = _parts(v); u3x, u3y = _parts(w)
    w1x, w1y = _parts(wx); w3x, w3y = _parts(wz)
    o1x = w1x - (-k3 * u2y); o1y = w1y - k3 * u2x
    o2x = -(k3 * u1y - k1 * u3y); o2y = k3 * u1x - k1 * u3x
    o3x = -w3x - (k1 * u2y); o3y = -w3y + k1 * u2x
    return _mk(G, o1x, o1y), _mk(G, o2x, o2y), _mk(G, o3x, o3y)


def to_physical(G: Grid, f):
    """fftBackwardTranspose (fft.c:54-78): [i][k][j] -> real [j][x][z], un-normalised C2R."""
    a = np.ascontiguousarray(np.transpose(f, (2, 0, 1)))
    return (np.fft.irfft2(a, s=(G.nx, G.nzp), axes=(1, 2), norm="forward")).astype(G.real)


def to_spectral(G: Grid, p):
    """fftForwardTranspose (fft.c:80-105): real [j][x][z] -> [i][k][j], un-normalised R2C."""
    a = np.fft.rfft2(p.astype(G.real), axes=(1, 2)).astype(G.cplx)
    return np.ascontiguousarray(np.transpose(a, (1, 2, 0)))


def rotor(G: Grid, u, v, w, ox, oy, oz):
    """rotorkernel (convolution_kernels.cu:69-148) on physical arrays: m = (u/N2) x (omega/N2)."""
    R = G.real; n2 = R(G.N2)
    u = u / n2; v = v / n2; w = w / n2; ox = ox / n2; oy = oy / n2; oz = oz / n2
    return (v * oz - w * oy).astype(R), (w * ox - u * oz).astype(R), (u * oy - v * ox).astype(R)


def convolution(G: Grid, u, v, w, return_physical=False):
    """convolution (convolution.c:3-53) -> (H1, H2, H3) [+ physical u, v, w]."""
    wz = deriv_y(G, u)
    wx = deriv_y(G, w)
    dudy_walls = (wz[0, 0, 0], wz[0, 0, G.ny - 1])          # readUtau (convolution.c:14-16)
    ox, oy, oz = calc_omega(G, wx, wz, u, v, w)
    pu, pv, pw = (to_physical(G, a) for a in (u, v, w))
    pox, poy, poz = (to_physical(G, a) for a in (ox, oy, oz))
    m1, m2, m3 = rotor(G, pu, pv, pw, pox, poy, poz)
    H = tuple(dealias(G, to_spectral(G, m)) for m in (m1, m2, m3))
    if return_physical:
        return H, (pu, pv, pw), dudy_walls
    return H


def calc_hvg(G: Grid, h1, h2, h3):
    """calcHvg (nonLinear_kernels.cu:94-190, 212-232) -> (h_v, h_g)."""
    k1 = G.k1[:, None, None]; k3 = G.k3[None, :, None]; kk = G.kk[:, :, None]
    h1x, h1y = _parts(h1); h3x, h3y = _parts(h3)
    hvx = -(-k1 * h1y - k3 * h3y); hvy = -k1 * h1x - k3 * h3x
    hgx = -(k3 * h1y - k1 * h3y); hgy = k3 * h1x - k1 * h3x
    hv = deriv_y(G, _mk(G, hvx, hvy))
    hg = _mk(G, hgx, hgy)
    ax, ay = _parts(hv); bx, by = _parts(h2)
    return _mk(G, ax - kk * bx, ay - kk * by), hg


def impose_symmetry(G: Grid, f):
    """imposeSymetry (imposeSymetry.c:5-18): C2R, R2C, /N2."""
    s = to_spectral(G, to_physical(G, f))
    return (s / G.real(G.N2)).astype(G.cplx)      # normalizekernel (dealias.cu:72-97): u /= int N2


def rk_step_1(G: Grid, u, u_yy, r, dt, n):
    """rk_step_1 (RK3_kernels.cu:6-93); u_yy = deriv_YY(u) (RKstep_1 :155-177)."""
    R = G.real
    nu = R(1.0) / G.REYNOLDS
    kk = G.kk[:, :, None]; dt = R(dt)
    al = R(ALPHA[n]); ds = R(DSEDA[n])
    ux, uy = _parts(u); wx, wy = _parts(u_yy); rx, ry = _parts(r)
    ox = ux + dt * (al * nu * (wx - kk * ux) + ds * rx)
    oy = uy + dt * (al * nu * (wy - kk * uy) + ds * ry)
    return _mk(G, ox, oy)


def rk_step_2(G: Grid, u_w, r, dt, n):
    """rk_step_2 (RK3_kernels.cu:95-153)."""
    R = G.real
    c = R(dt) * R(GAMMAD[n])
    return (u_w + c * r).astype(G.cplx)


# ---------------------------------------------------------------------------
# mean flow (host fp32 in the reference; state-type here)
# ---------------------------------------------------------------------------
class MeanFlow:
    """meanUevol.c: (kx,kz)=(0,0) ODE.  u, N, uw are the .x parts of u_host, N_host, uw_host."""

    def __init__(self, G: Grid):
        self.G = G
        R = G.real
        self.u = np.zeros(G.ny, R); self.N = np.zeros(G.ny, R); self.uw = np.zeros(G.ny, R)
        self.utau = (R(0), R(0))
        ny = G.ny; j = np.arange(ny)
        y = G.ymesh_h
        self.a = (y(j + 1) - y(j)).astype(R)
        self.b = (y(j - 1) - y(j)).astype(R)

    @staticmethod
    def _tri(x, a, b, c):
        """solve_tridiagonal_in_place_destructive (meanUevol.c:86-114), state-type."""
        R = x.dtype.type
        n = x.size
        c = c.copy(); x = x.copy()
        c[0] = c[0] / b[0]; x[0] = x[0] / b[0]
        for i in range(1, n):
            m = R(1.0 / F64(b[i] - a[i] * c[i - 1]))
            c[i] = c[i] * m
            x[i] = (x[i] - a[i] * x[i - 1]) * m
        for i in range(n - 2, -1, -1):
            x[i] = x[i] - c[i] * x[i + 1]
        return x

    def _w2(self):
        """second-derivative compact weights in the state type (meanUevol.c:228-236, 275-280)."""
        a, b = self.a, self.b
        R = self.G.real
        den = a * a * a - b * b * b - R(4.0) * a * a * b + R(4.0) * b * b * a
        A = R(-12.0) * b / den; B = R(12.0) * a / den; C = -A - B
        den2 = a * a * a - R(4.0) * a * a * b + R(4.0) * a * b * b - b * b * b
        al = -(-b * b * b - a * b * b + a * a * b) / den2
        be = -(a * a * a + b * a * a - b * b * a) / den2
        return A, B, C, al, be

    def second_derivative(self, u):
        """secondDerivative (meanUevol.c:223-345)."""
        G = self.G; R = G.real; ny = G.ny
        A, B, C, al, be = self._w2()
        y = G.ymesh_h
        up = al.copy(); cd = np.ones(ny, R); lo = be.copy()
        lo[0] = R(0); up[ny - 1] = R(0)
        a0 = R(y(1) - y(0)); b0 = R(y(2) - y(0))
        up[0] = (a0 + b0) / (R(2.0) * a0 - b0)
        aN = R(y(ny - 2) - y(ny - 1)); bN = R(y(ny - 3) - y(ny - 1))
        lo[ny - 1] = (aN + bN) / (R(2.0) * aN - bN)
        aux = np.zeros(ny, R)
        aux[1:-1] = A[1:-1] * u[2:] + C[1:-1] * u[1:-1] + B[1:-1] * u[:-2]
        A0 = R(6.0) / ((a0 - b0) * (R(2.0) * a0 - b0)); B0 = R(-6.0) * a0 / ((a0 * b0 - b0 * b0) * (R(2.0) * a0 - b0)); E0 = -A0 - B0
        aux[0] = E0 * u[0] + A0 * u[1] + B0 * u[2]
        AN = R(6.0) / ((aN - bN) * (R(2.0) * aN - bN)); BN = R(-6.0) * aN / ((aN * bN - bN * bN) * (R(2.0) * aN - bN)); EN = -AN - BN
        aux[ny - 1] = EN * u[ny - 1] + AN * u[ny - 2] + BN * u[ny - 3]
        return self._tri(aux, lo, cd, up)

    def implicit_step(self, u, betha_rk, dt):
        """implicitStepMeanU (meanUevol.c:347-437)."""
        G = self.G; R = G.real; ny = G.ny
        A, B, C, al, be = self._w2()
        nu = R(1.0) / G.REYNOLDS
        f = nu * R(dt) * R(betha_rk)
        lo = be - f * B; cd = R(1.0) - f * C; up = al - f * A
        lo[0] = 0; cd[0] = 1; up[0] = 0; lo[1] = 0
        lo[ny - 1] = 0; cd[ny - 1] = 1; up[ny - 1] = 0; up[ny - 2] = 0
        aux = np.zeros(ny, R)
        aux[1:-1] = al[1:-1] * u[2:] + u[1:-1] + be[1:-1] * u[:-2]
        return self._tri(aux, lo, cd, up)

    def forcing(self):
        """forcing (meanUevol.c:201-221): constant mass flux; the integral starts at j=10 (Q12)."""
        G = self.G; R = G.real; ny = G.ny
        n2 = R(G.N2)
        y = G.ymesh_h
        Qt = R(0.0)
        for j in range(10, ny - 1):
            dy = R(y(j + 1) - y(j))
            Qt = R(Qt + R(0.5) * (self.u[j + 1] + self.u[j]) * dy)
        self.u[1:ny - 1] = self.u[1:ny - 1] + (G.QVELOCITY * n2 - Qt) / G.LY

    def step_1(self, n):
        """meanURKstep_1 (meanUevol.c:439-464)."""
        G = self.G; R = G.real
        d2 = self.second_derivative(self.u.copy())
        self.uw = ((R(ALPHA[n]) + R(BETHA[n])) * (R(1.0) / G.REYNOLDS) * d2 + R(DSEDA[n]) * self.N).astype(R)

    def step_2(self, dt, n):
        """meanURKstep_2 (meanUevol.c:466-487)."""
        R = self.G.real
        self.uw = (R(dt) * (self.uw + R(GAMMAD[n]) * self.N)).astype(R)
        self.uw = self.implicit_step(self.uw, BETHA[n], dt)
        self.u = (self.u + self.uw).astype(R)
        self.forcing()

    def read_utau(self, dudy0, dudyN):
        """readUtau (meanUevol.c:178-199)."""
        R = self.G.real; n2 = R(self.G.N2)
        f = lambda d: R(np.sqrt(R(1.0) / self.G.REYNOLDS * np.abs(R(d.real) / n2)))
        self.utau = (f(dudy0), f(dudyN))


# ---------------------------------------------------------------------------
# L4 time integrator
# ---------------------------------------------------------------------------
def calc_dt(G: Grid, umax, vmax, wmax, cfl=0.5):
    """calcDt (RK3.c:64-109), state-type arithmetic; maxima over valid points (Q7)."""
    R = G.real
    n2 = R(G.N2); pi = R(np.arccos(F32(-1.0)))
    nx = R(G.nx); nz = R(G.nz); CFL = R(cfl)
    with np.errstate(divide="ignore"):
        c = (nx / (R(2.0) * pi * G.LX)) * np.abs(R(umax) / n2) + np.abs(R(vmax) / n2) / G.DELTA_Y + (nx / (R(2.0) * pi * G.LZ)) * np.abs(R(wmax) / n2)
        dt_c = CFL / R(c)
    third = R(1.0) / R(3.0)
    dt_v = CFL * G.REYNOLDS / (R(1.0) / (G.DELTA_Y * G.DELTA_Y) + (third * nx / G.LX) * (third * nx / G.LX) + (third * nz / G.LZ) * (third * nz / G.LZ))
    return R(min(dt_c, dt_v)), R(dt_c), R(dt_v)


def statistics(G: Grid, u, v, w):
    """calcSt (statistics.cu:161-243) + sumElementsComplex (fft.c:239-269): the four
    per-plane sums (-<uv> weights, u, v, w energy) before the sqrt of the printout."""
    R = G.real; n2 = R(G.N2)
    wgt = np.ones((1, G.nz, 1), R); wgt[0, 0, 0] = R(0.5)

    def ssum(ax, ay):                       # reals 0 .. 2NZ-3 of every row (k <= NZ-2)
        return (ax[:, :G.nz - 1, :].sum(axis=(0, 1), dtype=F64) + ay[:, :G.nz - 1, :].sum(axis=(0, 1), dtype=F64)).astype(R)

    ux, uy = u.real / n2, u.imag / n2
    vx, vy = v.real / n2, v.imag / n2
    wx, wy = w.real / n2, w.imag / n2
    out = [ssum(R(-2.0) * vx * ux * wgt, R(-2.0) * vy * uy * wgt)]
    for ax, ay in ((ux, uy), (vx, vy), (wx, wy)):
        out.append(ssum(R(2.0) * ax * ax * wgt, R(2.0) * ay * ay * wgt))
    return out


class State:
    """Everything RKstep (RK3.c:111-193) keeps between substeps."""

    def __init__(self, G: Grid, ddv, g, umean=None):
        self.G = G
        z = lambda: np.zeros((G.nx, G.nz, G.ny), G.cplx)
        self.ddv = ddv.astype(G.cplx); self.g = g.astype(G.cplx)
        self.v = z(); self.dv = z(); self.R_ddv = z(); self.R_g = z()
        self.mean = MeanFlow(G)
        if umean is not None:
            self.mean.u = np.asarray(umean).real.astype(G.real)
        self.dt = G.real(0.0)
        self.last_dt_c = None
        self.umax = (0.0, 0.0, 0.0)
        self.counter = 0

    def init(self):
        """RK3.c:120-123."""
        G = self.G
        self.ddv = dealias(G, self.ddv); self.g = dealias(G, self.g)
        self.ddv = impose_symmetry(G, self.ddv); self.g = impose_symmetry(G, self.g)
        self.v, self.dv = calc_vdv(G, self.ddv)

    def calc_nl(self, n):
        """calcNL (nonLinear.c:3-34) -> (R_ddv, R_g) = (h_v, h_g); updates mean.N, umax."""
        G = self.G
        u, w = calc_uw(G, self.dv, self.g)
        if G.kx_int[0] == 0:                              # writeUmean (meanUevol.c:117-125)
            u[0, 0, :] = self.mean.u.astype(G.cplx)
        (h1, h2, h3), (pu, pv, pw), walls = convolution(G, u, self.v, w, return_physical=True)
        self.mean.read_utau(*walls)
        self.mean.N = h1[0, 0, :].real.astype(G.real)     # readNmean (meanUevol.c:127-133)
        self.umax = (np.abs(pu).max(), np.abs(pv).max(), np.abs(pw).max())
        return calc_hvg(G, h1, h2, h3)

    def substep(self, n, fixed_dt=None):
        """body of the n_step loop (RK3.c:129-167)."""
        G = self.G; dt = self.dt
        ddv_w = rk_step_1(G, self.ddv, deriv_yy(G, self.ddv), self.R_ddv, dt, n)
        g_w = rk_step_1(G, self.g, deriv_yy(G, self.g), self.R_g, dt, n)
        self.mean.step_1(n)
        self.R_ddv, self.R_g = self.calc_nl(n)
        if n == 0:
            d, dc, dv_ = calc_dt(G, *self.umax)
            self.last_dt_c = d
            self.dt_2 = G.real(fixed_dt) if (fixed_dt is not None and fixed_dt > 0) else d
        self.mean.step_2(dt, n)
        ddv_w = rk_step_2(G, ddv_w, self.R_ddv, dt, n)
        g_w = rk_step_2(G, g_w, self.R_g, dt, n)
        g_w = implicit(G, g_w, BETHA[n], dt)
        ddv_w, self.v, self.dv = bilaplacian(G, ddv_w, BETHA[n], dt)
        self.ddv = dealias(G, ddv_w); self.g = dealias(G, g_w)
        self.v = dealias(G, self.v); self.dv = dealias(G, self.dv)

    def step(self, fixed_dt=None):
        for n in range(3):
            self.substep(n, fixed_dt)
        if self.counter % 1000 == 0:                      # RK3.c:174-176
            self.ddv = impose_symmetry(self.G, self.ddv); self.g = impose_symmetry(self.G, self.g)
        self.dt = self.dt_2                               # one-step lag (RK3.c:179)
        self.counter += 1

    def run(self, nsteps, fixed_dt=None):
        """RKstep (RK3.c:111-193) for nsteps steps."""
        self.init()
        for _ in range(nsteps):
            self.step(fixed_dt)
        return self
