"""Multi-GPU parity check, launched with torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tests/mgpu_check.py

Every rank steps its kx slab / y-plane slab of the golden cases and compares the slab it owns with
the dumps of the reference build (tests/golden).  Exit code 0 = parity on every rank."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import channel_gpu_b200 as cg
from channel_gpu_b200 import nccl_boot
from tests.util import Golden, rel


def main():
    rank = int(os.environ["RANK"]); size = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
    dist.init_process_group("gloo")
    torch.cuda.set_device(local)
    comm = nccl_boot.init_comm(rank, size, local)
    worst = 0.0
    for case, sub, n, tol in (("64x21x9_f64", "steps2", 2, 1e-11), ("192x19x13_f32", "steps2", 2, 1e-4), ("64x21x9_f32", "steps3_freedt", 3, 5e-5)):
        g = Golden(case)
        ddv, gg, um = g.inputs()
        ch = cg.Channel(g.nx, g.ny, g.nz, precision=4 if g.prec == 32 else 8, device=local, rank=rank, size=size, nccl_comm=comm)
        if os.environ.get("CHAN_PEER", "1") != "0":
            ch.enable_peer_exchange()        # bulk NVLink peer copies; CHAN_PEER=0 keeps the NCCL send/recv exchange
        ch.RKstep(ddv, gg, n, fixed_dt=None if "freedt" in sub else g.dt, umean=um)
        Kx = 2 * (g.nx // 3) + 1
        s0, c0 = cg.partition(Kx, size, rank)
        ikx = np.arange(s0, s0 + c0); kx = np.where(ikx <= g.nx // 3, ikx, ikx - Kx); ii = np.where(kx >= 0, kx, kx + g.nx)
        for name in ("ddv", "g", "v", "dv", "R_ddv", "R_g"):
            full = g.full(f"{sub}/{name}")
            got = ch.get_field(name)[ii]; want = full[ii]
            # slab error against the norm of the WHOLE field: a slab of high wavenumbers alone carries
            # little energy, and fp32 round-off of the nonlinear term would dominate a per-slab ratio
            e = float(np.linalg.norm((got - want).ravel()) / np.linalg.norm(full.ravel())); worst = max(worst, e / tol)
            if e >= tol:
                print(f"[rank {rank}] {case} {name}: {e:.2e} >= {tol}", flush=True); sys.exit(1)
        sc = ch.scalars()
        ref_dt = g.raw(f"{sub}/dt")[0]
        if abs(sc["dt_c"] - ref_dt) > 2e-6 * ref_dt:
            print(f"[rank {rank}] {case} dt {sc['dt_c']} vs {ref_dt}", flush=True); sys.exit(1)
        if rank == 0:
            e = rel(ch.get_umean().real, g.raw(f"{sub}/umean").real)
            if e > 1e-6:
                print(f"[rank 0] {case} umean {e}", flush=True); sys.exit(1)
        st = ch.statistics()
        if not np.isfinite(st).all():
            sys.exit(1)
        # operator-level convolution on a multi-GPU handle (inputs sit in the local pencil fields, whatever exchange the
        # substep uses) against the reference build's dump; it must leave the step scalars alone
        if g.has("ops/conv_wx"):
            before = ch.scalars()
            H = ch.convolution(g.full("ops/uw_u"), g.full("ops/v0"), g.full("ops/uw_w"))
            otol = 1e-12 if g.prec == 64 else 1e-6
            for hh, nm in zip(H, ("conv_wx", "conv_wy", "conv_wz")):
                full = g.full("ops/" + nm)
                e = float(np.linalg.norm((hh[ii] - full[ii]).ravel()) / np.linalg.norm(full.ravel())); worst = max(worst, e / otol)
                if e >= otol:
                    print(f"[rank {rank}] {case} convolution {nm}: {e:.2e} >= {otol}", flush=True); sys.exit(1)
            after = ch.scalars()
            if after["umax"] != before["umax"] or after["utau"] != before["utau"] or after["dt_next"] != before["dt_next"]:
                print(f"[rank {rank}] {case} chan_convolution changed the step scalars", flush=True); sys.exit(1)
        ch.close()
    dist.barrier()
    print(f"[rank {rank}/{size}] multi-GPU parity OK (worst error/tolerance {worst:.2e})", flush=True)
    nccl_boot.destroy_comm(comm)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
