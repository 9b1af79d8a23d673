"""Creates the ncclComm_t the library needs for multi-GPU runs (chan_config.nccl_comm).

torch.distributed is the plumbing: rank 0 makes an NCCL unique id, the process group broadcasts
its 128 bytes, every rank calls ncclCommInitRank through ctypes on the libnccl.so.2 that torch has
already loaded (so the process uses one NCCL).  A C host would do the same with its own launcher."""
import ctypes as C


class _UniqueId(C.Structure):
    _fields_ = [("internal", C.c_byte * 128)]


_nccl = None


def _lib():
    global _nccl
    if _nccl is None:
        import torch  # noqa: F401  (loads the bundled libnccl.so.2 first)
        _nccl = C.CDLL("libnccl.so.2")
        _nccl.ncclGetUniqueId.argtypes = [C.POINTER(_UniqueId)]
        _nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, _UniqueId, C.c_int]
        _nccl.ncclCommDestroy.argtypes = [C.c_void_p]
        _nccl.ncclGetErrorString.restype = C.c_char_p
    return _nccl


def init_comm(rank, size, device):
    """Returns an ncclComm_t (as int) spanning the ranks of the default torch process group."""
    import torch
    import torch.distributed as dist
    L = _lib()
    torch.cuda.set_device(device)
    uid = _UniqueId()
    if rank == 0:
        rc = L.ncclGetUniqueId(C.byref(uid))
        if rc:
            raise RuntimeError("ncclGetUniqueId: %s" % L.ncclGetErrorString(rc).decode())
    payload = [bytes(uid.internal) if rank == 0 else None]
    dist.broadcast_object_list(payload, src=0)
    C.memmove(C.byref(uid), payload[0], 128)
    comm = C.c_void_p()
    rc = L.ncclCommInitRank(C.byref(comm), size, uid, rank)
    if rc:
        raise RuntimeError("ncclCommInitRank: %s" % L.ncclGetErrorString(rc).decode())
    return comm.value


def destroy_comm(comm):
    if comm:
        _lib().ncclCommDestroy(comm)
