#!/usr/bin/env python
"""BASELINE.json configs[4]: the wall-normal stage (batched Helmholtz / bi-Laplacian / compact-derivative solves, fused
kernels pencil_pre and pencil_post) and the plane stage against NY and the mode count, on the GPUs of this launch.

    python scripts/sweep_r02.py [--out profiles/sweep_r02.json]          (1 GPU)
    torchrun ... scripts/sweep_r02.py --out ...                           (N GPUs: bench.py is launched per point)

Every point is one `bench.py --grid NX,NY,NZ --profile` run (5 steps after 3 warm-up steps, CUDA-event stage timers);
the y <-> xz transposes have no kernel of their own (plane-major layout on one GPU, peer stores from pencil_pre / xfft_fwd
on several), so their cost is inside those kernels.  Rows: ny, modes (active (kx,kz) pairs of the whole grid), kernel,
ms per substep, algorithmic GB/s, fraction of the measured HBM peak."""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NYS = (97, 129, 193, 257, 385, 513, 633)
GRIDS = ((96, 97), (192, 193), (384, 385), (768, 769))          # (NX, NZ): 4225, 16641, 66049, 263169 active modes


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "sweep_r02.json"))
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--variants", default="auto,chain,lanes")
    a = ap.parse_args()
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    env0 = {"auto": {}, "chain": {"CHAN_TUNE_LANES": "0", "CHAN_TUNE_PENCIL_SEG": "0"}, "lanes": {"CHAN_TUNE_LANES": "3"}}
    rows = []
    for ny in NYS:
        for nx, nz in GRIDS:
            modes = (2 * (nx // 3) + 1) * ((2 * nz - 2) // 3 + 1)
            if modes * ny * 16 * 20 > 120e9 * a.gpus:
                continue
            for var in a.variants.split(","):
                cmd = [sys.executable]
                if a.gpus > 1:
                    cmd += ["-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(a.gpus), "--master-addr", "127.0.0.1", "--master-port", "29688"]
                cmd += [os.path.join(ROOT, "bench.py"), "--gpus", str(a.gpus), "--grid", "%d,%d,%d" % (nx, ny, nz), "--steps", "5", "--warmup", "3", "--profile"]
                r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, env=dict(os.environ, **env0[var]), cwd=ROOT)
                line = [l for l in r.stdout.splitlines() if l.startswith("{")]
                if r.returncode or not line:
                    print("FAILED", ny, nx, nz, var, r.stderr[-300:], flush=True)
                    continue
                d = json.loads(line[-1])
                for k in ("pencil_pre", "pencil_post", "xfft_inv", "zstage", "xfft_fwd"):
                    s = d["stages"][k]
                    rows.append({"n_gpus": a.gpus, "ny": ny, "nx": nx, "nz": nz, "modes": modes, "variant": var, "kernel": k, "ms_per_substep": s["ms_per_substep"],
                                 "alg_GBps": s["GBps"], "frac_of_hbm_peak": (s["GBps"] or 0) / peak})
                print("ny=%d modes=%d %-5s step %.3f ms  pre %.3f post %.3f z %.3f" % (ny, modes, var, d["ms_per_step"], d["stages"]["pencil_pre"]["ms_per_substep"],
                      d["stages"]["pencil_post"]["ms_per_substep"], d["stages"]["zstage"]["ms_per_substep"]), flush=True)
                json.dump({"peak_GBps": peak, "how": __doc__, "rows": rows}, open(a.out, "w"), indent=0)


if __name__ == "__main__":
    main()
