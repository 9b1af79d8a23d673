"""CPU tests of the plain-C host side (host/): the run.conf reader with the reference's entry points
(config.c:4-42) and the text formats of the statistics files (statistics.cu:191-232)."""
import ctypes as C
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SAMPLE = b'''application:
{
	NX = 384;
	NY = 193;
	NZ = 193; #Points in physical space are 2*NZ-2
	input:
	{
 		G = "-";
		DDV = "-";
		UMEAN = "-";
		# G = "/somewhere/G.h5";
	};
	output:
	{
		G = "/tmp/out/G.01.h5";
		DDV = "/tmp/out/ddV.01.h5";
		UMEAN = "/tmp/out/Umean2.bin";
	};
	path = "/tmp/out/";
	REYNOLDS = 10000;  // optional run-time keys
	NSTEPS = 7; FIXED_DT = 0.0025; /* C comment */ PRECISION = 8;
};
'''


class Domain(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("nx", "ny", "nz", "rank", "size", "iglobal")]


class Paths(C.Structure):
    _fields_ = [(n, C.c_char * 100) for n in ("ginput", "goutput", "ddvinput", "ddvoutput", "umeaninput", "umeanoutput", "path")]


class Opts(C.Structure):
    _fields_ = [("reynolds", C.c_double), ("qvelocity", C.c_double), ("cfl", C.c_double), ("fixed_dt", C.c_double),
                ("nsteps", C.c_int), ("precision", C.c_int), ("checkpoint_every", C.c_int),
                ("checkpoint", C.c_char * 100), ("restart", C.c_char * 100)]


@pytest.fixture(scope="module")
def host():
    from channel_gpu_b200.build import build_lib
    build_lib()
    import importlib.util
    spec = importlib.util.spec_from_file_location("host_build", os.path.join(ROOT, "host", "build.py"))
    m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
    so, exe = m.build_host()
    assert os.path.exists(exe)
    return C.CDLL(so)


def test_run_conf_reference_syntax(host, tmp_path):
    cfg = C.create_string_buffer(64 * 1024)
    assert host.conf_parse_text(cfg, SAMPLE) == 0
    d = Domain(); host.read_domain_from_config(C.byref(d), cfg)
    assert (d.nx, d.ny, d.nz) == (384, 193, 193)
    p = Paths(); assert host.read_filenames_from_config(C.byref(p), cfg) == 0
    assert p.ginput == b"-" and p.ddvinput == b"-" and p.umeaninput == b"-"
    assert p.goutput == b"/tmp/out/G.01.h5" and p.umeanoutput == b"/tmp/out/Umean2.bin" and p.path == b"/tmp/out/"
    o = Opts(); host.read_run_opts_from_config(C.byref(o), cfg)
    assert (o.reynolds, o.nsteps, o.fixed_dt, o.precision, o.cfl, o.qvelocity) == (10000.0, 7, 0.0025, 8, 0.5, 1.8)
    # defaults = the reference's compile-time constants when the optional keys are absent
    assert host.conf_parse_text(cfg, b'application: { NX = 128; NY = 128; NZ = 65; };') == 0
    host.read_run_opts_from_config(C.byref(o), cfg)
    assert (o.reynolds, o.nsteps, o.precision) == (3250.0, 30000, 4)
    assert (o.checkpoint_every, o.checkpoint, o.restart) == (0, b"-", b"-")
    assert host.conf_parse_text(cfg, b'application: { NX = 8; checkpoint: { EVERY = 50; FILE = "/tmp/c.ckpt"; RESTART = "/tmp/old.ckpt"; }; };') == 0
    host.read_run_opts_from_config(C.byref(o), cfg)
    assert (o.checkpoint_every, o.checkpoint, o.restart) == (50, b"/tmp/c.ckpt", b"/tmp/old.ckpt")
    assert host.conf_parse_text(cfg, b'application: { NX = 128; NY = 128; NZ = 65; };') == 0
    assert host.read_filenames_from_config(C.byref(p), cfg) == 7       # all seven names missing
    # through a file, like main.c:34
    f = tmp_path / "run.conf"; f.write_bytes(SAMPLE)
    assert host.read_config_file(cfg, str(f).encode()) == 0
    assert host.read_config_file(cfg, str(tmp_path / "absent.conf").encode()) != 0


def test_run_conf_of_the_reference_itself(host):
    ref = "/root/reference/run.conf"
    if not os.path.exists(ref):
        pytest.skip("reference checkout not present")
    cfg = C.create_string_buffer(64 * 1024)
    assert host.read_config_file(cfg, ref.encode()) == 0
    d = Domain(); host.read_domain_from_config(C.byref(d), cfg)
    assert (d.nx, d.ny, d.nz) == (128, 128, 65)                          # run.conf:3-5
    p = Paths(); assert host.read_filenames_from_config(C.byref(p), cfg) == 0
    assert p.ginput == b"-" and p.path.endswith(b"/")


def test_run_conf_syntax_errors(host):
    cfg = C.create_string_buffer(64 * 1024)
    for bad in (b"application: { NX = ; };", b"application: { NX = 3;", b'a = "unterminated;', b"} ;", b"x 3;"):
        assert host.conf_parse_text(cfg, bad) != 0


def test_statistics_text_format(host, tmp_path):
    ny = 5
    sums = (C.c_double * (4 * ny))(*([0.0, -1.0, 4.0, 2.25, 0.0] + [3.2068] * ny + [0.25] * ny + [1.0] * ny))
    path = (str(tmp_path) + "/").encode()
    for _ in range(2):
        assert host.append_calcst_files(path, sums, ny, 4) == 0
    lines = open(tmp_path / "RSTRSS.dat").read().split("\n")
    assert lines[0] == " 0.000000 -nan 2.000000 1.500000 0.000000 " and lines[1] == lines[0] and lines[2] == ""
    assert open(tmp_path / "VRMS.dat").read().split("\n")[0] == " 0.500000" * ny + " "
    # golden: the reference's own URMS.dat line has the same shape (one leading blank per value, trailing blank)
    gold = open(os.path.join(ROOT, "tests", "golden", "64x21x9_f32_steps3_dat", "URMS.dat")).read()
    assert gold.startswith(" ") and gold.endswith(" \n") and len(gold.split()) == 21


class H5Info(C.Structure):
    _fields_ = [("dims", C.c_uint64 * 3), ("elem_bytes", C.c_int), ("data_address", C.c_uint64), ("data_bytes", C.c_uint64),
                ("dataset_header", C.c_uint64), ("eof_address", C.c_uint64)]


@pytest.mark.parametrize("prec", [4, 8])
def test_hdf5_field_container(host, tmp_path, prec):
    """The field files of the reference: one contiguous HDF5 dataset "u", float, {NX, NY, 2NZ} (hit_mpi.c:272-279).
    No libhdf5 / h5py exists here, so the file is checked against the byte positions the HDF5 File Format
    Specification fixes (superblock v0, object header v1, group B-tree + symbol-table node + local heap)."""
    import numpy as np
    import struct
    nx, ny, nz = 12, 9, 5
    rng = np.random.default_rng(7)
    cplx = np.complex64 if prec == 4 else np.complex128
    a = (rng.standard_normal((nx, nz, ny)) + 1j * rng.standard_normal((nx, nz, ny))).astype(cplx)     # bytes = [i][k][j][re,im]
    name = str(tmp_path / "G.h5").encode()
    assert host.write_field(name, a.ctypes.data, nx, ny, nz, prec) == 0
    raw = open(name, "rb").read()
    # superblock
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0 and raw[13] == 8 and raw[14] == 8
    leaf_k, node_k = struct.unpack_from("<HH", raw, 16); assert (leaf_k, node_k) == (4, 16)
    base, freesp, eof, drv = struct.unpack_from("<QQQQ", raw, 24)
    assert base == 0 and freesp == drv == 2 ** 64 - 1 and eof == len(raw)
    name_off, root_hdr, cache = struct.unpack_from("<QQI", raw, 56)
    btree, heap = struct.unpack_from("<QQ", raw, 80)
    assert name_off == 0 and cache == 1
    # root object header: version 1, one Symbol Table message pointing at the same B-tree and heap
    ver, _, nmsg, refc, hsize = struct.unpack_from("<BBHII", raw, root_hdr)
    assert (ver, nmsg, refc) == (1, 1, 1)
    mtype, msize = struct.unpack_from("<HH", raw, root_hdr + 16); assert (mtype, msize) == (0x11, 16)
    assert struct.unpack_from("<QQ", raw, root_hdr + 24) == (btree, heap)
    # group B-tree node -> symbol-table node -> link "u" through the local heap
    assert raw[btree:btree + 4] == b"TREE" and raw[btree + 4] == 0 and raw[btree + 5] == 0
    assert struct.unpack_from("<H", raw, btree + 6)[0] == 1
    snod = struct.unpack_from("<Q", raw, btree + 32)[0]
    assert raw[snod:snod + 4] == b"SNOD" and struct.unpack_from("<H", raw, snod + 6)[0] == 1
    link_off, dset = struct.unpack_from("<QQ", raw, snod + 8)
    assert raw[heap:heap + 4] == b"HEAP"
    hsz, hfree, hdata = struct.unpack_from("<QQQ", raw, heap + 8)
    assert raw[hdata + link_off:hdata + link_off + 2] == b"u\x00" and hfree + 16 <= hsz
    # dataset header: dataspace {NX, NY, 2NZ}, IEEE little-endian float of the state size, contiguous layout
    ver, _, nmsg = struct.unpack_from("<BBH", raw, dset); assert ver == 1
    pos = dset + 16; seen = {}
    for _ in range(nmsg):
        t, sz = struct.unpack_from("<HH", raw, pos); seen[t] = pos + 8; pos += 8 + sz; assert sz % 8 == 0
    sp = seen[1]; assert raw[sp] == 1 and raw[sp + 1] == 3 and struct.unpack_from("<QQQ", raw, sp + 8) == (nx, ny, 2 * nz)
    dtp = seen[3]; assert raw[dtp] == 0x11 and struct.unpack_from("<I", raw, dtp + 4)[0] == prec
    assert struct.unpack_from("<HH", raw, dtp + 8) == (0, 8 * prec)
    assert tuple(raw[dtp + 12:dtp + 16]) == ((23, 8, 0, 23) if prec == 4 else (52, 11, 0, 52))
    assert struct.unpack_from("<I", raw, dtp + 16)[0] == (127 if prec == 4 else 1023)
    lay = seen[8]; assert raw[lay] == 3 and raw[lay + 1] == 1
    addr, nbytes = struct.unpack_from("<QQ", raw, lay + 2)
    assert nbytes == a.nbytes and addr + nbytes == len(raw) and raw[addr:] == a.tobytes()
    # the reader follows the addresses and returns the same bytes; a wrong shape or type is refused
    info = H5Info(); assert host.h5min_info(name, C.byref(info)) == 0
    assert tuple(info.dims) == (nx, ny, 2 * nz) and info.elem_bytes == prec and info.data_address == addr
    b = np.empty_like(a); assert host.read_field(name, b.ctypes.data, nx, ny, nz, prec) == 0
    assert np.array_equal(a, b)
    assert host.read_field(name, b.ctypes.data, nx, ny, nz + 1, prec) != 0
    assert host.read_field(name, b.ctypes.data, nx, ny, nz, 12 - prec) != 0
    # a raw file name keeps the raw payload
    rawname = str(tmp_path / "G.bin").encode()
    assert host.write_field(rawname, a.ctypes.data, nx, ny, nz, prec) == 0 and open(rawname, "rb").read() == a.tobytes()
