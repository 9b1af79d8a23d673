"""CPU-side checks of the drop-in boundary: the shared library loads, exports every
symbol include/channel_b200.h declares, and its pure-host helpers behave."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as ge
    from channel_gpu_b200.build import build_lib
    build_lib()
    import channel_gpu_b200 as cg
    return cg


def test_every_declared_symbol_is_exported(lib):
    L = lib.load_library()
    hdr = open(os.path.join(ROOT, "include", "channel_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(chan_[a-z_0-9]+)\s*\(", hdr)))
    assert len(names) >= 35
    for n in names:
        assert hasattr(L, n), n


def test_partition_is_even_and_complete(lib):
    for n in (85, 97, 193, 257, 385, 633, 683, 1365):
        for size in (1, 2, 3, 4, 8):
            pos = 0
            counts = []
            for r in range(size):
                s, c = lib.partition(n, size, r)
                assert s == pos
                pos += c; counts.append(c)
            assert pos == n and max(counts) - min(counts) <= 1
    L = lib.load_library()
    assert L.chan_partition(10, 0, 0, None, None) != 0
    assert L.chan_active_kx(384) == 257 and L.chan_active_kz(193) == 129
    assert L.chan_active_kx(2048) == 1365 and L.chan_active_kz(769) == 513


def test_create_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(lib.ChannelError):
        lib.Channel(64, 21, 9)
