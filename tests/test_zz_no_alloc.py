"""chan_deriv_y_pencils_static — the entry point for hosts that cannot allocate.

A caller whose device memory is entirely owned by its own arena (or who must avoid the
device-wide stall of cudaMalloc) still needs wall-normal derivatives.  This entry point
keeps its Thomas scratch in a fixed array inside the library, so the call itself
allocates nothing.  The test drives it through the C ABI on a run of reference-layout
pencils and compares with chan_deriv_y (the handle-scratch path) on the same data,
twice in a row with different inputs: the second call finds the first call's leftovers
in the shared array and must not be affected by them.  It also checks the size limit.

Collected last on purpose (file name), so a run that stops at the first failure has
already exercised everything else.
"""
import numpy as np
import pytest

from oracle import kmm
from tests.util import Golden, rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("prec", [8, 4])
def test_static_scratch_variant_matches_handle_variant_twice(prec):
    import torch
    import channel_gpu_b200 as cg
    g = Golden("64x21x9_f64")
    nx, ny, nz = g.nx, g.ny, g.nz
    cplx = np.complex128 if prec == 8 else np.complex64
    G = kmm.Grid(nx, ny, nz, np.float64 if prec == 8 else np.float32)
    ch = cg.Channel(nx, ny, nz, precision=prec)
    tol = 1e-12 if prec == 8 else 1e-6          # the operator tolerances of tests/test_gpu_parity.py
    ddv, gg, _ = g.inputs()
    for field in (kmm.dealias(G, ddv.astype(cplx)), kmm.dealias(G, (3.0 * gg - 0.5 * ddv).astype(cplx))):
        want = ch.deriv_Y_HO_double(field)                       # existing entry point (chan_deriv_y)
        d = torch.from_numpy(np.ascontiguousarray(field)).cuda()
        ch.deriv_y_pencils_static(d, nx * nz)                     # all NX*NZ pencils of the reference layout
        got = d.cpu().numpy()
        assert rel(got, want) < tol
        if prec == 8:
            assert np.array_equal(got, want)                      # same tables, same operations per row
        assert rel(got, kmm.deriv_y(G, field)) < tol              # and the oracle
    # a sub-range of pencils, addressed by pointer offset into the same array
    field = kmm.dealias(G, ddv.astype(cplx))
    want = ch.deriv_Y_HO_double(field)
    d = torch.from_numpy(np.ascontiguousarray(field)).cuda()
    first, count = 5 * nz + 2, 3 * nz
    ch._ck(ch.L.chan_deriv_y_pencils_static(ch.h, d.data_ptr() + first * ny * d.element_size(), count), "static sub-range")
    got = d.cpu().numpy().reshape(nx * nz, ny)
    assert rel(got[first:first + count], want.reshape(nx * nz, ny)[first:first + count]) < tol
    assert np.array_equal(got[:first], field.reshape(nx * nz, ny)[:first])      # untouched outside the range
    # beyond the fixed array: refused with a message, nothing launched
    n0 = ch.launch_count()
    too_many = (2 << 20) // (2 * ny) + 1
    rc = ch.L.chan_deriv_y_pencils_static(ch.h, d.data_ptr(), too_many)
    assert rc == 1 and b"static array" in ch.L.chan_last_error(ch.h)
    assert ch.launch_count() == n0
    ch.close()
