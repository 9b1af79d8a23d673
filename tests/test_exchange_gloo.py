"""The N>1 path on the CPU: two (and three) gloo ranks perform the pencil <-> plane exchange of the
multi-GPU code with the library's own slab / layout arithmetic (chan_partition,
chan_exchange_layout — the functions comm.cu calls before every NCCL send/recv) on numpy buffers,
and check that (a) the plane side sees every active mode of its planes exactly where the FFT
kernels' SegMap addressing expects it, and (b) the way back restores the pencil slabs."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _code(j, ikx, kz):
    return j * 1_000_000.0 + ikx * 1_000.0 + kz


def _worker(rank, size, nx, ny, nz, port, q):
    import channel_gpu_b200 as cg
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        L = cg.load_library()
        Kx, Kz = L.chan_active_kx(nx), L.chan_active_kz(nz)
        kx0, kxl = cg.partition(Kx, size, rank)
        j0, nyl = cg.partition(ny, size, rank)
        M = kxl * Kz
        # pencil side: F[j][ikx_local][kz], all planes, my kx slab
        jj, ii, kk = np.meshgrid(np.arange(ny), np.arange(kx0, kx0 + kxl), np.arange(Kz), indexing="ij")
        pencil = _code(jj, ii, kk).reshape(-1).copy()
        plane = np.full(nyl * Kx * Kz, -1.0)
        lay = [cg.exchange_layout(nx, ny, nz, size, rank, p) for p in range(size)]
        # ---- to planes
        reqs = []; recv_t = {}
        for p in range(size):
            s = torch.from_numpy(pencil[lay[p]["send_off"]: lay[p]["send_off"] + lay[p]["send_cnt"]].copy())
            if p == rank:
                plane[lay[p]["recv_off"]: lay[p]["recv_off"] + lay[p]["recv_cnt"]] = s.numpy()
                continue
            recv_t[p] = torch.empty(lay[p]["recv_cnt"], dtype=torch.float64)
            reqs.append(dist.isend(s, p)); reqs.append(dist.irecv(recv_t[p], p))
        for r in reqs:
            r.wait()
        for p, t in recv_t.items():
            plane[lay[p]["recv_off"]: lay[p]["recv_off"] + lay[p]["recv_cnt"]] = t.numpy()
        # SegMap addressing of plane_core.cuh: base[s] + (jl*width[s] + ikx - start[s])*Kz + kz
        ok = True
        for jl in range(nyl):
            for ikx in range(Kx):
                s = max(p for p in range(size) if lay[p]["peer_kx0"] <= ikx)
                off = lay[s]["recv_off"] + (jl * lay[s]["peer_kxl"] + ikx - lay[s]["peer_kx0"]) * Kz
                ok &= bool(np.array_equal(plane[off:off + Kz], _code(j0 + jl, ikx, np.arange(Kz))))
        assert (plane >= 0).all()
        # ---- and back
        back = np.full_like(pencil, -1.0)
        reqs = []; recv_t = {}
        for p in range(size):
            s = torch.from_numpy(plane[lay[p]["recv_off"]: lay[p]["recv_off"] + lay[p]["recv_cnt"]].copy())
            if p == rank:
                back[lay[p]["send_off"]: lay[p]["send_off"] + lay[p]["send_cnt"]] = s.numpy()
                continue
            recv_t[p] = torch.empty(lay[p]["send_cnt"], dtype=torch.float64)
            reqs.append(dist.isend(s, p)); reqs.append(dist.irecv(recv_t[p], p))
        for r in reqs:
            r.wait()
        for p, t in recv_t.items():
            back[lay[p]["send_off"]: lay[p]["send_off"] + lay[p]["send_cnt"]] = t.numpy()
        ok &= bool(np.array_equal(back, pencil))
        # ---- the same exchange as ONE-SIDED copies (what the bulk NVLink peer copies of the GPU path do): every rank tells
        # its destinations where it writes (chan_peer_put_layout); the destinations only apply what they are told
        put = [cg.peer_put_layout(nx, ny, nz, size, rank, p) for p in range(size)]
        plane2 = np.full(nyl * Kx * Kz, -1.0); back2 = np.full_like(pencil, -1.0)
        for direction in ("out", "back"):
            src = pencil if direction == "out" else plane
            dst = plane2 if direction == "out" else back2
            mine = [(put[p][direction + "_dst"], src[put[p][direction + "_src"]: put[p][direction + "_src"] + put[p][direction + "_cnt"]].copy()) for p in range(size)]
            for p in range(size):          # all-to-all of (offset, payload) pairs through object gathers
                gathered = [None] * size
                dist.all_gather_object(gathered, mine[p])
                # gathered[r] = what rank r addressed to destination p; only destination p applies it
                if p == rank:
                    for off, data in gathered:
                        assert (dst[off: off + len(data)] < 0).all()          # nobody else wrote there
                        dst[off: off + len(data)] = data
        ok &= bool(np.array_equal(plane2, plane)) and bool(np.array_equal(back2, pencil))
        # a sub-range of planes (one plane batch) is the same triple shifted by rows * modes of the slab
        for p in range(size):
            pj0, pnyl = cg.partition(ny, size, p)
            if pnyl >= 2:
                r0, nr = 1, pnyl - 1
                seg = pencil[put[p]["out_src"] + r0 * M: put[p]["out_src"] + (r0 + nr) * M]
                ok &= bool(np.array_equal(seg.reshape(nr, kxl, Kz)[:, 0, 0], _code(pj0 + r0 + np.arange(nr), kx0, 0)))
        # every rank's slabs tile the active modes and the planes
        tot = torch.tensor([float(kxl), float(nyl)])
        dist.all_reduce(tot)
        ok &= (int(tot[0]) == Kx and int(tot[1]) == ny)
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("size,grid", [(2, (64, 21, 9)), (2, (48, 19, 13)), (3, (48, 19, 13))])
def test_pencil_plane_exchange_layout(size, grid):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + size * 7 + grid[0] % 13
    procs = [ctx.Process(target=_worker, args=(r, size, *grid, port, q)) for r in range(size)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
    assert sorted(r for r, _ in res) == list(range(size))
    assert all(ok for _, ok in res)
