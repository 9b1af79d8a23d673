#!/usr/bin/env python
"""Pack the raw dumps that tests/golden/run_golden.sh left in gpurun_out/golden/
into the committed fixtures tests/golden/<case>.npz (+ the reference's text
outputs, verbatim).  Run here after the gpurun call; the GPU box never needs it.

Complex fields keep the reference layout [i][k][j]; only the ACTIVE modes of the
2/3 rule (dealias.cu:37) are stored, as `<name>` with shape
[n_active_kx][n_active_kz][NY].  At pack time every dump is checked to be zero on
truncated modes up to FFT round-off (meta["leak"]), so nothing is lost.
"""
import os, sys, glob, json, shutil
import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(ROOT, "..", "..", "gpurun_out", "golden")


def pack(case):
    dims, prec = case.split("_f")
    nx, ny, nz = (int(a) for a in dims.split("x"))
    ct = np.complex64 if prec == "32" else np.complex128
    rt = np.float32 if prec == "32" else np.float64
    i = np.arange(nx); kx = np.where(i < nx // 2, i, i - nx)
    ax = np.abs(kx) <= nx // 3
    kzmax = (2 * nz - 2) // 3
    out = {}
    meta = {"nx": nx, "ny": ny, "nz": nz, "prec": int(prec), "dt": 0.005, "subs": []}
    for sub in sorted(os.listdir(os.path.join(SRC, case))):
        d = os.path.join(SRC, case, sub)
        if not os.path.isdir(d):
            continue
        meta["subs"].append(sub)
        for f in sorted(glob.glob(os.path.join(d, "*.bin"))):
            name = os.path.basename(f)[:-4]
            key = f"{sub}/{name}"
            if name == "dt":
                out[key] = np.fromfile(f, np.float64)
            elif name == "umean":
                out[key] = np.fromfile(f, ct)
            elif name == "phys":
                out[key] = np.fromfile(f, rt).reshape(ny, nx, 2 * nz)[:, :, :2 * nz - 2].copy()
            else:
                a = np.fromfile(f, ct).reshape(nx, nz, ny)
                act = a[ax][:, :kzmax + 1]
                rest = a.copy(); rest[np.ix_(ax, np.arange(nz) <= kzmax)] = 0
                # imposeSymetry (imposeSymetry.c:5-18) does not re-truncate, so dumps taken
                # before the next dealias carry FFT round-off on truncated modes
                leak = float(np.abs(rest).max() / max(np.abs(a).max(), 1e-300))
                assert leak < (1e-3 if prec == "32" else 1e-10), f"{key}: {leak} outside the active modes"
                meta.setdefault("leak", {})[key] = leak
                out[key] = act.copy()
    np.savez_compressed(os.path.join(ROOT, case + ".npz"), meta=json.dumps(meta), **out)
    print(case, "%.2f MB" % (os.path.getsize(os.path.join(ROOT, case + ".npz")) / 1e6))


if __name__ == "__main__":
    for case in sorted(os.listdir(SRC)):
        if os.path.isdir(os.path.join(SRC, case)):
            pack(case)
    # the reference's own text outputs of a 3-step run (formats of statistics.cu:173-232, meanUevol.c:511-557)
    dst = os.path.join(ROOT, "64x21x9_f32_steps3_dat")
    os.makedirs(dst, exist_ok=True)
    for f in glob.glob(os.path.join(SRC, "64x21x9_f32", "steps3", "*.dat")):
        shutil.copy(f, dst)
    shutil.copy(os.path.join(SRC, "ref_time.log"), os.path.join(ROOT, "ref_time_r01.log"))
