"""Which y-stage kernel is closer to the exact R_phi = d/dy[-i(k1 H1 + k3 H3)] - kk H2 ?  (GPU box)
One RK3 step with dt = 0 on a thin 24 x NY x 13 grid; H1..3 are read back and the same expression is evaluated on the
host in extended precision (numpy longdouble) with the oracle's coefficients; the thread-per-chain and the on-chip
kernels are both compared with it."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import kmm
from bench import synthetic_fields

LD = np.longdouble


def exact_rphi(G, H1, H2, H3):
    c = G.coef(); ny = G.ny
    k1 = G.k1d[:, None, None].astype(LD); k3 = G.k3d[None, :, None].astype(LD)
    hv = -1j * (k1 * H1.astype(np.clongdouble) + k3 * H3.astype(np.clongdouble))
    A1, B1, C1 = (c[k].astype(LD) for k in ("A1", "B1", "C1")); al1, be1 = c["al1"].astype(LD), c["be1"].astype(LD)
    A0, B0, C0, al0 = (LD(v) for v in c["w0"]); AN, BN, CN, alN = (LD(v) for v in c["wN"])
    rhs = np.zeros_like(hv)
    rhs[..., 1:-1] = A1[1:-1] * hv[..., 2:] + C1[1:-1] * hv[..., 1:-1] + B1[1:-1] * hv[..., :-2]
    rhs[..., 0] = A0 * hv[..., 0] + B0 * hv[..., 1] + C0 * hv[..., 2]
    rhs[..., -1] = AN * hv[..., -1] + BN * hv[..., -2] + CN * hv[..., -3]
    l = be1.copy(); u = al1.copy(); l[0] = 0; u[0] = al0; l[-1] = alN; u[-1] = 0
    cp = np.zeros(ny, LD); dp = np.zeros_like(rhs)
    cp[0] = u[0]; dp[..., 0] = rhs[..., 0]
    for j in range(1, ny):
        m = LD(1) / (LD(1) - l[j] * cp[j - 1]); cp[j] = u[j] * m
        dp[..., j] = (rhs[..., j] - l[j] * dp[..., j - 1]) * m
    x = np.zeros_like(rhs); x[..., -1] = dp[..., -1]
    for j in range(ny - 2, -1, -1):
        x[..., j] = dp[..., j] - cp[j] * x[..., j + 1]
    kk = (k1 * k1 + k3 * k3)
    return x - kk * H2.astype(np.clongdouble)


def main():
    import channel_gpu_b200 as cg
    nx, nz = 24, 13
    for ny in (97, 193, 257, 385, 633):
        ddv, gg, um = synthetic_fields(nx, ny, nz, np.complex128)
        G = kmm.Grid(nx, ny, nz, np.float64)
        out = {}
        for name, env in (("chain", {"CHAN_TUNE_LANES": "0", "CHAN_TUNE_PENCIL_SEG": "0", "CHAN_TUNE_POST_DEEP": "0"}), ("lanes", {"CHAN_TUNE_LANES": "1", "CHAN_TUNE_PENCIL_SEG": "0"})):
            os.environ.update(env)
            ch = cg.Channel(nx, ny, nz, precision=8)
            ch.RKstep(ddv, gg, 1, fixed_dt=1e-3, umean=um)
            out[name] = {k: ch.get_field(k) for k in ("R_ddv", "wx", "wy", "wz", "dv", "v")}
            ch.close()
        # H of the last substep differs between the runs at round-off (v, dv of the earlier substeps do): each kernel is
        # compared with the exact expression of ITS OWN H
        e = {}
        for k in ("chain", "lanes"):
            ex = exact_rphi(G, *[out[k][f] for f in ("wx", "wy", "wz")]) * G.active[:, :, None]
            n = float(np.linalg.norm(ex.astype(np.complex128).ravel()))
            e[k] = float(np.linalg.norm((out[k]["R_ddv"].astype(np.clongdouble) - ex).astype(np.complex128).ravel())) / n
            hv = np.linalg.norm((ex + (G.kkd[:, :, None] * out[k]["wy"])).astype(np.complex128).ravel())
        d = float(np.linalg.norm((out["chain"]["R_ddv"] - out["lanes"]["R_ddv"]).ravel())) / n
        dH = max(float(np.linalg.norm((out["chain"][f] - out["lanes"][f]).ravel()) / np.linalg.norm(out["chain"][f].ravel())) for f in ("wx", "wy", "wz"))
        print("NY=%d  R_phi error vs extended precision: chain %.2e  lanes %.2e   chain-lanes %.2e  (H chain-lanes %.1e, |hv'|/|R| %.1e)" % (ny, e["chain"], e["lanes"], d, dH, hv / n), flush=True)


if __name__ == "__main__":
    main()
