#!/usr/bin/env python
"""Seeded input fields for the golden runs of the reference build.

    python tests/golden/make_inputs.py NX NY NZ OUTDIR [32|64]

Writes OUTDIR/{ddv,g,umean}.bin as raw complex arrays in the reference layout
[i][k][j] (RK3_kernels.cu:16-18) in the state type of the build (float2 or
double2).  The double files hold the float32 values widened, so both builds see
bit-identical inputs.  tests/golden/run_golden.sh feeds these to
oracle/_ref/ref_* on the GPU box; tests load the same arrays through
`golden_inputs()` below, so nothing has to be regenerated at test time.

Fields are smooth random data on ALL modes (the dealias / symmetry operators are
part of what is checked): phi = a(k) * (xi1 (1-y^2) + xi2 y (1-y^2) + xi3 (1-y^2)^2),
g likewise, with a(k) a Gaussian in (kx, kz) times N2 (spectral arrays carry the
un-normalised factor N2 = NX*(2NZ-2), convolution_kernels.cu:93), plus a laminar
mean profile 1.35 (1-y^2) N2.
"""
import sys
import numpy as np

SEED = 20261106


def mesh(ny):
    # channel.h:21,50-51 ; derivatives_nu_double.cu:60  (float product, double tanh)
    dy = np.float32(2.0) / np.float32(ny - 1)
    j = np.arange(ny, dtype=np.float32)
    x = (j * dy).astype(np.float64) - 1.0
    return np.tanh(2.0 * x) / np.tanh(2.0)


def golden_inputs(nx, ny, nz, amp=0.35):
    rng = np.random.default_rng(SEED + 7 * nx + 3 * ny + nz)
    n2 = nx * (2 * nz - 2)
    y = mesh(ny)
    kx = np.where(np.arange(nx) < nx // 2, np.arange(nx), np.arange(nx) - nx).astype(np.float64)
    kz = np.arange(nz, dtype=np.float64)
    a = np.exp(-(kx[:, None] / (nx / 7.0)) ** 2 - (kz[None, :] / ((2 * nz - 2) / 7.0)) ** 2)
    shapes = np.stack([1 - y ** 2, y * (1 - y ** 2), (1 - y ** 2) ** 2])       # [3][ny]

    def field(scale):
        xi = rng.standard_normal((nx, nz, 3)) + 1j * rng.standard_normal((nx, nz, 3))
        f = np.einsum("xzs,sj->xzj", xi, shapes) * a[:, :, None] * (scale * n2)
        return f.astype(np.complex64)

    ddv = field(amp * 6.0)
    g = field(amp * 2.0)
    umean = (1.35 * (1 - y ** 2) * n2).astype(np.float32).astype(np.complex64)
    return ddv, g, umean


def main():
    nx, ny, nz = (int(a) for a in sys.argv[1:4])
    out = sys.argv[4]
    prec = int(sys.argv[5]) if len(sys.argv) > 5 else 32
    import os
    os.makedirs(out, exist_ok=True)
    ddv, g, um = golden_inputs(nx, ny, nz)
    ct = np.complex64 if prec == 32 else np.complex128
    ddv.astype(ct).tofile(os.path.join(out, "ddv.bin"))
    g.astype(ct).tofile(os.path.join(out, "g.bin"))
    um.astype(ct).tofile(os.path.join(out, "umean.bin"))


if __name__ == "__main__":
    main()
