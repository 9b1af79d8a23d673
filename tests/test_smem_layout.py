"""Design check of the z-stage shared-memory row layouts (channel_gpu_b200/csrc/plane_core.cuh, launch_z in
kernels_plane.cu) with the bank-conflict model of scripts/smem_model.py: for the transform shapes of the
BASELINE grids every phase but the Hermitian unpack must be conflict-free with the packed layouts, and the
padded layout they replaced must be measurably worse (ncu: 131 M wavefronts per launch for 74 M ideal)."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
import smem_model


@pytest.mark.parametrize("n,radices,xp,nthr,kz", [(384, (16, 6, 4), 1, 128, 129), (768, (16, 12, 4), 1, 192, 257), (128, (8, 4, 4), 3, 128, 43)])
def test_packed_layouts_are_conflict_free(n, radices, xp, nthr, kz):
    rep = smem_model.report(n, radices, xp, nthr, kz)
    for label, (got, ideal) in rep["packed (new)"]:
        if label != "unpack loads":
            assert got == ideal, label
    tot_new = sum(r[1][0] for r in rep["packed (new)"]); tot_old = sum(r[1][0] for r in rep["padded (old)"])
    ideal = sum(r[1][1] for r in rep["packed (new)"])
    assert tot_new <= 1.08 * ideal
    assert tot_old >= 1.2 * ideal
