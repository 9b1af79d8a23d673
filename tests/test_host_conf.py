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
                ("nsteps", C.c_int), ("precision", C.c_int)]


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
