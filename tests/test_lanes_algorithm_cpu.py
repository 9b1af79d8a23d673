"""The on-chip wall-normal stage (channel_gpu_b200/csrc/kernels_pencil_lanes.cu) solves tridiagonal systems with the lanes of a
warp ALONG y: each lane sweeps S rows with 0 entering, the per-lane affine maps (product of a, last value) are combined by a
Kogge-Stone scan that stops as soon as the neglected products fall below 2^-54, and each lane adds (running product) x
(entering value).  This is a numpy restatement of exactly that procedure (same segment length, same depth rule, same order of
the two sweeps) checked against the sequential Thomas algorithm of the oracle on the reference's matrices: the compact d/dy
matrix (derivatives_nu_double.cu:227-274), the Helmholtz matrix A - kk M for smooth and rough modes
(hemholzt_nu_double.cu:136-185) and the implicit matrix M - D (A - kk M) (implicitStep_nu_double.cu:133-163), NY = 97 .. 633.
It pins the truncation rule without a GPU; the kernels themselves are compared with the thread-per-chain kernels and the
reference build in the -m gpu tests."""
import numpy as np
import pytest

from oracle import kmm

S = 4          # rows per lane (lanes_S)
WARP = 32


def scan_depth(amax):
    """depth rule of scan_maps: lb = 1022 - exponent field of amax (amax < 2^-lb); first depth with lb << depth >= 54"""
    if amax <= 0.0:
        return 0
    expo = (np.float64(amax).view(np.uint64) >> np.uint64(52)) & np.uint64(0x7ff)
    lb = 1022 - int(expo)
    if lb <= 0:
        return 5
    d = 0
    while (lb << d) < 54 and d < 5:
        d += 1
    return d


def recurrence_lanes(a, b, reverse=False):
    """x_j = b_j + a_j x_(j-1) (or from the other end), the way the lanes kernel evaluates it.  a: [n], b: [n, nrhs]."""
    n = len(a)
    if reverse:
        return recurrence_lanes(a[::-1], b[::-1], False)[0][::-1], None
    nl = (n + S - 1) // S
    W = (nl + WARP - 1) // WARP
    L = W * WARP
    ap = np.zeros(L * S); bp = np.zeros((L * S, b.shape[1]))
    ap[:n] = a; bp[:n] = b
    ap = ap.reshape(L, S); bp = bp.reshape(L, S, -1)
    y = np.zeros_like(bp); P = np.zeros((L, S))
    prev = np.zeros((L, b.shape[1])); p = np.ones(L)
    for r in range(S):                                    # every lane: local sweep with 0 entering
        prev = bp[:, r] + ap[:, r, None] * prev
        y[:, r] = prev
        p = p * ap[:, r]; P[:, r] = p
    A = p.copy(); Y = prev.copy()
    depths = []
    carry = np.zeros(b.shape[1])
    xin = np.zeros((L, b.shape[1]))
    for w in range(W):                                    # one warp: truncated Kogge-Stone, then the carry of the warps before
        sl = slice(w * WARP, (w + 1) * WARP)
        Aw = A[sl].copy(); Yw = Y[sl].copy()
        depth = scan_depth(np.abs(Aw).max())
        depths.append(depth)
        for st in range(depth):
            off = 1 << st
            Ap = np.roll(Aw, off); Yp = np.roll(Yw, off, axis=0)
            on = np.arange(WARP) >= off
            Yw = np.where(on[:, None], Yw + Aw[:, None] * Yp, Yw)
            Aw = np.where(on, Aw * Ap, Aw)
        Ae = np.concatenate([[1.0], Aw[:-1]]); Ye = np.concatenate([np.zeros((1, b.shape[1])), Yw[:-1]])
        xin[sl] = Ye + Ae[:, None] * carry
        carry = Yw[-1] + Aw[-1] * carry
    x = y + P[:, :, None] * xin[:, None, :]
    return x.reshape(L * S, -1)[:n], depths


def solve_lanes(lo, di, up, rhs):
    """Thomas with precomputed pivots, both sweeps as lane recurrences (var_lanes / d1_lanes)."""
    n = len(di)
    piv = np.zeros(n); cp = np.zeros(n)
    c = 0.0
    for j in range(n):                                    # pivot_lanes_kernel: the only serial part, once per matrix
        piv[j] = 1.0 / (di[j] - lo[j] * c); c = up[j] * piv[j]; cp[j] = c
    d, dep_f = recurrence_lanes(-lo * piv, rhs * piv[:, None])
    x, _ = recurrence_lanes(-cp, d, reverse=True)
    return x, dep_f


def rel(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


@pytest.mark.parametrize("ny", [97, 193, 385, 633])
def test_lane_scans_equal_thomas_on_the_reference_matrices(ny):
    G = kmm.Grid(24, ny, 13, np.float64)
    c = G.coef()
    rng = np.random.default_rng(ny)
    y = c["y"]
    f = np.stack([np.cos(3 * np.pi * y) * (1 - y ** 2), rng.standard_normal(ny) * (1 - y ** 2) ** 2, np.sin(11 * y) + y ** 3, rng.standard_normal(ny)], axis=1)
    # ---- compact d/dy: beta f'_(j-1) + f'_j + alpha f'_(j+1) = rhs, one-sided closures at the walls
    l = c["be1"].copy(); u = c["al1"].copy(); d = np.ones(ny)
    l[0] = 0.0; u[0] = c["w0"][3]; l[-1] = c["wN"][3]; u[-1] = 0.0
    want = kmm.thomas(l, d, u, f.T.astype(np.complex128)).real.T
    got, depth = solve_lanes(l, d, u, f)
    assert rel(got, want) < 5e-15
    assert max(depth) <= 4                                  # |product of 4 rows| ~ 0.27^4: three or four doubling steps, never the full five
    # ---- Helmholtz (A - kk M) and implicit (M - D(A - kk M)) matrices, Dirichlet rows at the walls
    A2, B2, C2, al2, be2 = c["A2"], c["B2"], c["C2"], c["al2"], c["be2"]
    for kk, D in ((0.0, None), (1.0, None), (40.0 ** 2 + 4 * 30.0 ** 2, None), (1.0, 2e-7), (9000.0, 6e-8), (4.0, 5e-6)):
        if D is None:
            lo = B2 - kk * be2; di = C2 - kk; up = A2 - kk * al2
        else:
            lo = be2 - D * B2 + D * kk * be2; di = 1.0 - D * C2 + D * kk; up = al2 - D * A2 + D * kk * al2
        lo = lo.copy(); di = di.copy(); up = up.copy()
        lo[0] = up[0] = lo[-1] = up[-1] = 0.0; di[0] = di[-1] = 1.0; lo[1] = 0.0; up[-2] = 0.0
        rhs = al2[:, None] * np.roll(f, -1, axis=0) + f + be2[:, None] * np.roll(f, 1, axis=0)
        rhs[0] = rhs[-1] = 0.0
        want = kmm.thomas(lo, di, up, rhs.T.astype(np.complex128)).real.T
        got, depth = solve_lanes(lo, di, up, rhs)
        assert rel(got, want) < 2e-13, (kk, D, rel(got, want))
        if D is None and kk <= 1.0:
            assert max(depth) == 5                          # smooth Helmholtz modes couple the whole pencil: full scan
        if D is not None and D <= 2e-7:
            assert max(depth) <= 4                          # the implicit matrix is close to the mass matrix: never the full scan


def test_depth_rule():
    assert scan_depth(0.0) == 0 and scan_depth(1.0) == 5 and scan_depth(0.9) == 5
    assert scan_depth(2.0 ** -54) == 0 or scan_depth(2.0 ** -55) == 0
    for a in (0.5, 0.27 ** 4, 1e-3, 1e-9, 3e-17):
        k = scan_depth(a)
        assert k == 5 or a ** (1 << k) < 2.0 ** -53          # what the scan leaves out is below half an ulp
