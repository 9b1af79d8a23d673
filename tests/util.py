"""Shared helpers for the tests: golden fixtures and error norms."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    """relative L2 error of a against b"""
    a = np.asarray(a).astype(np.complex128).ravel()
    b = np.asarray(b).astype(np.complex128).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))


class Golden:
    """One packed case of tests/golden/pack_golden.py (dumps of the reference build)."""

    def __init__(self, case):
        z = np.load(os.path.join(GOLDEN, case + ".npz"))
        self.meta = json.loads(str(z["meta"]))
        self.z = z
        m = self.meta
        self.nx, self.ny, self.nz = m["nx"], m["ny"], m["nz"]
        self.prec = m["prec"]
        self.real = np.float32 if self.prec == 32 else np.float64
        self.cplx = np.complex64 if self.prec == 32 else np.complex128
        self.dt = m["dt"]
        i = np.arange(self.nx)
        kx = np.where(i < self.nx // 2, i, i - self.nx)
        self.ax = np.abs(kx) <= self.nx // 3
        self.kzmax = (2 * self.nz - 2) // 3

    def has(self, key):
        return key in self.z.files

    def raw(self, key):
        return self.z[key]

    def full(self, key):
        """complex field expanded to the reference layout [NX][NZ][NY]"""
        a = self.z[key]
        out = np.zeros((self.nx, self.nz, self.ny), self.cplx)
        out[np.ix_(self.ax, np.arange(self.nz) <= self.kzmax)] = a
        return out

    def inputs(self):
        from tests.golden.make_inputs import golden_inputs
        ddv, g, um = golden_inputs(self.nx, self.ny, self.nz)
        return ddv.astype(self.cplx), g.astype(self.cplx), um.astype(self.cplx)
