"""ctypes binding of include/channel_b200.h plus a thin object mirroring the
reference's driver interface (RK3.c / channel.h names in the docstrings).

Host arrays are numpy complex64/complex128 in the reference layout [NX][NZ][NY]
(RK3_kernels.cu:16-18).  Operator-level calls take device memory; when given
numpy arrays they stage them through torch CUDA tensors (torch is used for
device memory only).
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

FIELDS = {"ddv": 0, "g": 1, "v": 2, "dv": 3, "R_ddv": 4, "R_g": 5, "u": 6, "w": 7, "wx": 8, "wy": 9, "wz": 10}


class ChannelError(RuntimeError):
    pass


class _Config(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int), ("precision", C.c_int),
                ("reynolds", C.c_double), ("qvelocity", C.c_double), ("cfl", C.c_double),
                ("rank", C.c_int), ("size", C.c_int), ("device", C.c_int), ("plane_batch", C.c_int),
                ("nccl_comm", C.c_void_p)]


class _Scalars(C.Structure):
    _fields_ = [("dt", C.c_double), ("dt_next", C.c_double), ("dt_c", C.c_double), ("dt_v", C.c_double),
                ("umax", C.c_double * 3), ("utau", C.c_double * 2), ("counter", C.c_int), ("time", C.c_double)]


def lib_path():
    return os.path.join(_HERE, "lib", "libchannel_b200.so")


def load_library():
    """Loads the CUDA library; raises if it has not been built (no fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    p = lib_path()
    if not os.path.exists(p):
        raise ChannelError("%s is missing: run `python -m channel_gpu_b200.build` (nvcc, sm_100a). "
                           "There is no CPU fallback." % p)
    L = C.CDLL(p)
    vp, ci, cd = C.c_void_p, C.c_int, C.c_double
    sig = {
        "chan_config_default": [C.POINTER(_Config)],
        "chan_create": [C.POINTER(_Config), C.POINTER(vp)],
        "chan_set_fields": [vp, vp, vp], "chan_get_fields": [vp, vp, vp],
        "chan_get_field": [vp, ci, vp], "chan_set_field": [vp, ci, vp],
        "chan_get_field_slab": [vp, ci, vp], "chan_set_field_slab": [vp, ci, vp],
        "chan_get_field_dev": [vp, ci, vp], "chan_set_field_dev": [vp, ci, vp],
        "chan_get_fields_dev": [vp, vp, vp], "chan_set_fields_dev": [vp, vp, vp],
        "chan_checkpoint_save": [vp, vp], "chan_checkpoint_load": [vp, vp],
        "chan_set_umean": [vp, vp], "chan_get_umean": [vp, vp], "chan_get_nmean": [vp, vp],
        "chan_get_scalars": [vp, C.POINTER(_Scalars)],
        "chan_set_fixed_dt": [vp, cd], "chan_set_dt": [vp, cd],
        "chan_rk_init": [vp], "chan_rk_step": [vp, ci], "chan_rk_substep": [vp, ci], "chan_rk_end_step": [vp],
        "chan_sync": [vp], "chan_statistics": [vp, vp],
        "chan_set_timing": [vp, ci], "chan_get_stage_ms": [vp, C.POINTER(cd), C.POINTER(C.c_long)],
        "chan_dealias": [vp, vp], "chan_deriv_y": [vp, vp], "chan_deriv_yy": [vp, vp], "chan_helmholtz": [vp, vp],
        "chan_implicit": [vp, vp, cd, cd], "chan_bilaplacian": [vp, vp, vp, vp, cd, cd],
        "chan_calc_uw": [vp, vp, vp, vp, vp], "chan_convolution": [vp, vp, vp, vp, vp, vp, vp],
        "chan_calc_hvg": [vp, vp, vp, vp], "chan_impose_symmetry": [vp, vp, vp],
        "chan_fft_backward_transpose": [vp, vp, vp],
        "chan_deriv_y_pencils_static": [vp, vp, C.c_long],
        "chan_ipc_export": [vp, vp], "chan_ipc_import": [vp, ci, vp],
        "chan_partition": [ci, ci, ci, C.POINTER(ci), C.POINTER(ci)],
        "chan_active_kx": [ci], "chan_active_kz": [ci],
        "chan_exchange_layout": [ci, ci, ci, ci, ci, ci, C.POINTER(C.c_long)],
        "chan_peer_put_layout": [ci, ci, ci, ci, ci, ci, C.POINTER(C.c_long)],
    }
    for name, args in sig.items():
        f = getattr(L, name)
        f.argtypes = args
        f.restype = ci
    L.chan_destroy.argtypes = [vp]; L.chan_destroy.restype = None
    L.chan_last_error.argtypes = [vp]; L.chan_last_error.restype = C.c_char_p
    L.chan_launch_count.argtypes = [vp]; L.chan_launch_count.restype = C.c_long
    L.chan_checkpoint_bytes.argtypes = [vp]; L.chan_checkpoint_bytes.restype = C.c_size_t
    _LIB = L
    return L


def partition(n, size, rank):
    """chan_partition: even split of n items (active kx, y planes) over ranks."""
    L = load_library()
    s, c = C.c_int(), C.c_int()
    if L.chan_partition(n, size, rank, C.byref(s), C.byref(c)):
        raise ChannelError("bad partition arguments")
    return s.value, c.value


def exchange_layout(nx, ny, nz, size, rank, peer):
    """chan_exchange_layout as a dict (complex-element offsets/counts of one field's exchange with `peer`)."""
    L = load_library()
    out = (C.c_long * 8)()
    if L.chan_exchange_layout(nx, ny, nz, size, rank, peer, out):
        raise ChannelError("bad exchange_layout arguments")
    k = ("send_off", "send_cnt", "recv_off", "recv_cnt", "peer_kx0", "peer_kxl", "peer_j0", "peer_nyl")
    return dict(zip(k, [int(v) for v in out]))


def peer_put_layout(nx, ny, nz, size, rank, dest):
    """chan_peer_put_layout as a dict: the one-sided copies rank -> dest of one field, both directions."""
    L = load_library()
    out = (C.c_long * 6)()
    if L.chan_peer_put_layout(nx, ny, nz, size, rank, dest, out):
        raise ChannelError("bad peer_put_layout arguments")
    k = ("out_src", "out_dst", "out_cnt", "back_src", "back_dst", "back_cnt")
    return dict(zip(k, [int(v) for v in out]))


class Channel:
    """One GPU's time-stepper.  setUp + setRK3 + setRKmean (setUp.c:3, RK3.c:17, meanUevol.c:43)."""

    def __init__(self, nx, ny, nz, precision=8, reynolds=3250.0, qvelocity=1.8, cfl=0.5, device=-1, plane_batch=0,
                 rank=0, size=1, nccl_comm=None):
        self.L = load_library()
        cfg = _Config()
        self.L.chan_config_default(C.byref(cfg))
        cfg.nx, cfg.ny, cfg.nz, cfg.precision = nx, ny, nz, precision
        cfg.reynolds, cfg.qvelocity, cfg.cfl = reynolds, qvelocity, cfl
        cfg.device, cfg.plane_batch = device, plane_batch
        cfg.rank, cfg.size, cfg.nccl_comm = rank, size, nccl_comm
        self.rank, self.size = rank, size
        self.nx, self.ny, self.nz, self.precision = nx, ny, nz, precision
        self.cplx = np.complex64 if precision == 4 else np.complex128
        self.real = np.float32 if precision == 4 else np.float64
        h = C.c_void_p()
        rc = self.L.chan_create(C.byref(cfg), C.byref(h))
        self.h = h
        if rc:
            msg = self.L.chan_last_error(h).decode() if h else "allocation failed"
            if h:
                self.L.chan_destroy(h)
            self.h = None
            raise ChannelError("chan_create: %s" % msg)

    def close(self):
        if getattr(self, "h", None):
            self.L.chan_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc, what):
        if rc:
            raise ChannelError("%s failed (%d): %s" % (what, rc, self.L.chan_last_error(self.h).decode()))

    def _host(self, a):
        a = np.ascontiguousarray(a, dtype=self.cplx)
        if a.shape != (self.nx, self.nz, self.ny):
            raise ChannelError("expected shape %s, got %s" % ((self.nx, self.nz, self.ny), a.shape))
        return a

    # ---- state -----------------------------------------------------------------
    def set_fields(self, ddv, g):
        """readData / genRandData + H2D (io.c:3-34, 74-98)."""
        ddv, g = self._host(ddv), self._host(g)
        self._ck(self.L.chan_set_fields(self.h, ddv.ctypes.data, g.ctypes.data), "chan_set_fields")

    def get_field(self, name, out=None):
        """`out`: optional destination (e.g. a view of pinned memory) of the reference shape and dtype."""
        if out is None:      # several ranks: a rank fills only the rows of its own kx slab
            out = (np.empty if self.size == 1 else np.zeros)((self.nx, self.nz, self.ny), self.cplx)
        elif out.shape != (self.nx, self.nz, self.ny) or out.dtype != self.cplx or not out.flags["C_CONTIGUOUS"]:
            raise ChannelError("out must be a C-contiguous %s array of shape %s" % (np.dtype(self.cplx), (self.nx, self.nz, self.ny)))
        self._ck(self.L.chan_get_field(self.h, FIELDS[name], out.ctypes.data), "chan_get_field")
        return out

    def set_field(self, name, a):
        a = self._host(a)
        self._ck(self.L.chan_set_field(self.h, FIELDS[name], a.ctypes.data), "chan_set_field")

    def slab(self):
        """(first active-kx index, count) of this rank's pencil-stage slab and the reference row i of each slab row."""
        Kx = 2 * (self.nx // 3) + 1
        s0, c0 = partition(Kx, self.size, self.rank)
        ikx = np.arange(s0, s0 + c0); kx = np.where(ikx <= self.nx // 3, ikx, ikx - Kx)
        return s0, c0, np.where(kx >= 0, kx, kx + self.nx)

    def set_field_slab(self, name, a):
        """`a`: this rank's slab [Kxl][NZ][NY] in active-kx order (see slab())."""
        a = np.ascontiguousarray(a, dtype=self.cplx)
        if a.shape != (self.slab()[1], self.nz, self.ny):
            raise ChannelError("expected slab shape %s, got %s" % ((self.slab()[1], self.nz, self.ny), a.shape))
        self._ck(self.L.chan_set_field_slab(self.h, FIELDS[name], a.ctypes.data), "chan_set_field_slab")

    def get_field_slab(self, name, out=None):
        shp = (self.slab()[1], self.nz, self.ny)
        if out is None:
            out = np.empty(shp, self.cplx)
        elif out.shape != shp or out.dtype != self.cplx or not out.flags["C_CONTIGUOUS"]:
            raise ChannelError("out must be a C-contiguous %s array of shape %s" % (np.dtype(self.cplx), shp))
        self._ck(self.L.chan_get_field_slab(self.h, FIELDS[name], out.ctypes.data), "chan_get_field_slab")
        return out

    def checkpoint_save(self):
        """One rank's complete state as bytes (chan_checkpoint_save)."""
        buf = np.empty(int(self.L.chan_checkpoint_bytes(self.h)), np.uint8)
        self._ck(self.L.chan_checkpoint_save(self.h, buf.ctypes.data), "chan_checkpoint_save")
        return buf

    def checkpoint_load(self, buf):
        buf = np.ascontiguousarray(buf, dtype=np.uint8)
        self._ck(self.L.chan_checkpoint_load(self.h, buf.ctypes.data), "chan_checkpoint_load")

    def get_fields(self, out=None):
        """D2H + writeData payload (io.c:37-72).  `out`: optional (ddv, g) destination arrays."""
        return self.get_field("ddv", None if out is None else out[0]), self.get_field("g", None if out is None else out[1])

    def set_umean(self, um):
        """readU payload (meanUevol.c:166-176)."""
        um = np.ascontiguousarray(um, dtype=self.cplx)
        self._ck(self.L.chan_set_umean(self.h, um.ctypes.data), "chan_set_umean")

    def get_umean(self):
        out = np.empty(self.ny, self.cplx)
        self._ck(self.L.chan_get_umean(self.h, out.ctypes.data), "chan_get_umean")
        return out

    def get_nmean(self):
        out = np.empty(self.ny, np.float64)
        self._ck(self.L.chan_get_nmean(self.h, out.ctypes.data), "chan_get_nmean")
        return out

    def scalars(self):
        s = _Scalars()
        self._ck(self.L.chan_get_scalars(self.h, C.byref(s)), "chan_get_scalars")
        return dict(dt=s.dt, dt_next=s.dt_next, dt_c=s.dt_c, dt_v=s.dt_v, umax=tuple(s.umax), utau=tuple(s.utau), counter=s.counter, time=s.time)

    def set_fixed_dt(self, dt):
        self._ck(self.L.chan_set_fixed_dt(self.h, float(dt)), "chan_set_fixed_dt")

    def set_dt(self, dt):
        self._ck(self.L.chan_set_dt(self.h, float(dt)), "chan_set_dt")

    # ---- hot path -----------------------------------------------------------------
    def rk_init(self):
        """RKstep prologue (RK3.c:120-123)."""
        self._ck(self.L.chan_rk_init(self.h), "chan_rk_init")

    def rk_step(self, nsteps=1):
        """Body of RKstep's loop (RK3.c:127-191); asynchronous, see sync()."""
        self._ck(self.L.chan_rk_step(self.h, int(nsteps)), "chan_rk_step")

    def rk_substep(self, n):
        self._ck(self.L.chan_rk_substep(self.h, int(n)), "chan_rk_substep")

    def rk_end_step(self):
        self._ck(self.L.chan_rk_end_step(self.h), "chan_rk_end_step")

    def RKstep(self, ddv, g, nsteps, fixed_dt=None, umean=None):
        """RKstep(ddv, g, ...) (RK3.c:111-193) on host arrays: returns (ddv, g) after nsteps."""
        self.set_fields(ddv, g)
        if umean is not None:
            self.set_umean(umean)
        if fixed_dt is not None:
            self.set_fixed_dt(fixed_dt)
        self.rk_init()
        self.rk_step(nsteps)
        self.sync()
        return self.get_fields()

    def enable_peer_exchange(self):
        """Multi-GPU: map every other rank's exchange buffers (CUDA IPC) so that the two exchanges of a substep are bulk
        peer copies straight into them over NVLink.  torch.distributed carries the handles."""
        import torch.distributed as dist
        mine = C.create_string_buffer(9 * 64)
        self._ck(self.L.chan_ipc_export(self.h, mine), "chan_ipc_export")
        allh = [None] * self.size
        dist.all_gather_object(allh, bytes(mine.raw))
        for r, b in enumerate(allh):
            self._ck(self.L.chan_ipc_import(self.h, r, b), "chan_ipc_import")
        dist.barrier()

    def sync(self):
        self._ck(self.L.chan_sync(self.h), "chan_sync")

    def set_timing(self, on=True):
        self._ck(self.L.chan_set_timing(self.h, int(bool(on))), "chan_set_timing")

    def stage_ms(self):
        """Accumulated device milliseconds and counts per stage (pre, planes, scalar, post, whole step, xfft_inv, zstage, xfft_fwd, peer copies out, peer copies back) since the last call."""
        ms = (C.c_double * 10)(); cnt = (C.c_long * 10)()
        self._ck(self.L.chan_get_stage_ms(self.h, ms, cnt), "chan_get_stage_ms")
        return list(ms), list(cnt)

    def launch_count(self):
        return int(self.L.chan_launch_count(self.h))

    def statistics(self):
        """calcSt sums (statistics.cu:161-243): rows = (-<uv>, u, v, w)."""
        out = np.empty((4, self.ny), np.float64)
        self._ck(self.L.chan_statistics(self.h, out.ctypes.data), "chan_statistics")
        return out

    # ---- operator level (device pointers; numpy in/out staged through torch) -------------
    def _dev(self, a):
        import torch
        return torch.from_numpy(self._host(a)).cuda()

    def _op1(self, fn, name, u, *extra):
        d = self._dev(u)
        self._ck(fn(self.h, d.data_ptr(), *extra), name)
        return d.cpu().numpy()

    def dealias(self, u):
        return self._op1(self.L.chan_dealias, "chan_dealias", u)

    def deriv_Y_HO_double(self, u):
        return self._op1(self.L.chan_deriv_y, "chan_deriv_y", u)

    def deriv_YY_HO_double(self, u):
        return self._op1(self.L.chan_deriv_yy, "chan_deriv_yy", u)

    def hemholztSolver_double(self, u):
        return self._op1(self.L.chan_helmholtz, "chan_helmholtz", u)

    def implicitSolver_double(self, u, betha, dt):
        return self._op1(self.L.chan_implicit, "chan_implicit", u, float(betha), float(dt))

    def bilaplaSolver_double(self, ddv, betha, dt):
        import torch
        a = self._dev(ddv); b = torch.zeros_like(a); c = torch.zeros_like(a)
        self._ck(self.L.chan_bilaplacian(self.h, a.data_ptr(), b.data_ptr(), c.data_ptr(), float(betha), float(dt)), "chan_bilaplacian")
        return a.cpu().numpy(), b.cpu().numpy(), c.cpu().numpy()

    def calcUW(self, f, g):
        import torch
        fd, gd = self._dev(f), self._dev(g)
        u = torch.zeros_like(fd); w = torch.zeros_like(fd)
        self._ck(self.L.chan_calc_uw(self.h, u.data_ptr(), w.data_ptr(), fd.data_ptr(), gd.data_ptr()), "chan_calc_uw")
        return u.cpu().numpy(), w.cpu().numpy()

    def convolution(self, ux, uy, uz):
        import torch
        a, b, c = self._dev(ux), self._dev(uy), self._dev(uz)
        wx = torch.zeros_like(a); wy = torch.zeros_like(a); wz = torch.zeros_like(a)
        self._ck(self.L.chan_convolution(self.h, a.data_ptr(), b.data_ptr(), c.data_ptr(), wx.data_ptr(), wy.data_ptr(), wz.data_ptr()), "chan_convolution")
        return wx.cpu().numpy(), wy.cpu().numpy(), wz.cpu().numpy()

    def calcHvg(self, nlx, nly, nlz):
        a, b, c = self._dev(nlx), self._dev(nly), self._dev(nlz)
        self._ck(self.L.chan_calc_hvg(self.h, a.data_ptr(), b.data_ptr(), c.data_ptr()), "chan_calc_hvg")
        return a.cpu().numpy(), c.cpu().numpy()

    def imposeSymetry(self, u, v):
        a, b = self._dev(u), self._dev(v)
        self._ck(self.L.chan_impose_symmetry(self.h, a.data_ptr(), b.data_ptr()), "chan_impose_symmetry")
        return a.cpu().numpy(), b.cpu().numpy()

    def deriv_y_pencils_static(self, dev_tensor, n_pencils):
        """chan_deriv_y_pencils_static on a torch CUDA tensor holding [n_pencils][NY] complex, in place.
        Allocation-free inside the library (its scratch is a fixed array in the library image)."""
        self._ck(self.L.chan_deriv_y_pencils_static(self.h, dev_tensor.data_ptr(), int(n_pencils)), "chan_deriv_y_pencils_static")
        return dev_tensor

    def fftBackwardTranspose(self, u):
        """-> real [NY][NX][2NZ-2] (pad columns dropped)."""
        import torch
        a = self._dev(u)
        out = torch.zeros((self.ny, self.nx, 2 * self.nz), dtype=torch.float32 if self.precision == 4 else torch.float64, device="cuda")
        self._ck(self.L.chan_fft_backward_transpose(self.h, a.data_ptr(), out.data_ptr()), "chan_fft_backward_transpose")
        return out.cpu().numpy()[:, :, :2 * self.nz - 2]
