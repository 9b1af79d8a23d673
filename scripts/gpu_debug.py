"""First-contact script: every C-ABI entry point against the goldens / oracle, printing errors."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import kmm
from tests.util import Golden, rel
import channel_gpu_b200 as cg

def run(case):
    g = Golden(case)
    G = kmm.Grid(g.nx, g.ny, g.nz, g.real)
    ch = cg.Channel(g.nx, g.ny, g.nz, precision=4 if g.prec == 32 else 8)
    print("==", case)
    if g.has("ops/sym_ddv"):
        ddv, gg, um = g.inputs()
        P = g.full("ops/sym_ddv"); Gg = g.full("ops/sym_g")
        b0 = float(kmm.BETHA[0]); dt = g.dt
        print(" dealias", rel(ch.dealias(ddv), g.full("ops/dealias")))
        a, b = ch.imposeSymetry(kmm.dealias(G, ddv), kmm.dealias(G, gg))
        print(" sym", rel(a, P), rel(b, Gg))
        print(" derivY", rel(ch.deriv_Y_HO_double(P), g.full("ops/derivY")))
        print(" derivYY", rel(ch.deriv_YY_HO_double(P), g.full("ops/derivYY")))
        print(" helm", rel(ch.hemholztSolver_double(P), g.full("ops/helm")))
        print(" implicit", rel(ch.implicitSolver_double(P, b0, dt), g.full("ops/implicit")))
        a, b, c = ch.bilaplaSolver_double(P, b0, dt)
        print(" bil", rel(a, g.full("ops/bil_ddv")), rel(b, g.full("ops/bil_v")), rel(c, g.full("ops/bil_dv")))
        u, w = ch.calcUW(g.full("ops/dv0"), Gg)
        print(" uw", rel(u, g.full("ops/uw_u")), rel(w, g.full("ops/uw_w")))
        print(" phys", rel(ch.fftBackwardTranspose(P), g.raw("ops/phys")))
        H = ch.convolution(g.full("ops/uw_u"), g.full("ops/v0"), g.full("ops/uw_w"))
        print(" conv", [rel(h, g.full("ops/" + n)) for h, n in zip(H, ("conv_wx", "conv_wy", "conv_wz"))])
        hv, hg = ch.calcHvg(g.full("ops/conv_wx"), g.full("ops/conv_wy"), g.full("ops/conv_wz"))
        print(" hvg", rel(hv, g.full("ops/hvg_hv")), rel(hg, g.full("ops/hvg_hg")))
    for sub in g.meta["subs"]:
        if not sub.startswith("steps"):
            continue
        n = int(sub[5]); fixed = None if "freedt" in sub else g.dt
        ddv, gg, um = g.inputs()
        ch2 = cg.Channel(g.nx, g.ny, g.nz, precision=4 if g.prec == 32 else 8)
        ch2.RKstep(ddv, gg, n, fixed_dt=fixed, umean=um)
        S = kmm.State(G, ddv, gg, um).run(n, fixed)
        names = ("ddv", "g", "v", "dv", "R_ddv", "R_g")
        print(" %s vs ref   :" % sub, " ".join("%s %.1e" % (k, rel(ch2.get_field(k), g.full(f"{sub}/{k}"))) for k in names),
              "umean %.1e" % rel(ch2.get_umean().real, g.raw(f"{sub}/umean").real))
        print(" %s vs oracle:" % sub, " ".join("%s %.1e" % (k, rel(ch2.get_field(k), getattr(S, k))) for k in names),
              "umean %.1e" % rel(ch2.get_umean().real, S.mean.u))
        sc = ch2.scalars()
        print("   dt_c %.9e ref %.9e | utau %s oracle %s" % (sc["dt_c"], g.raw(f"{sub}/dt")[0], sc["utau"], S.mean.utau))
        ch2.close()
    ch.close()

for case in sys.argv[1:] or ["64x21x9_f64", "64x21x9_f32", "192x19x13_f32"]:
    run(case)
