"""Runs the plane-stage FFT phases (fft_core.cuh / plane_core.cuh) on the CPU: the same
__host__ __device__ code the kernels execute, with the threads of a block emulated in a loop,
against O(N^2) DFTs (tests/host/fft_host_test.cu).  No GPU needed."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc not available")
def test_fft_phases_on_host(tmp_path):
    exe = str(tmp_path / "fft_host_test")
    subprocess.run(["nvcc", "-O1", "-std=c++17", "-w", "-o", exe,
                    os.path.join(ROOT, "tests", "host", "fft_host_test.cu"),
                    os.path.join(ROOT, "channel_gpu_b200", "csrc", "kernels_tables.cu")], check=True, timeout=900)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:]
    assert "WORST" in r.stdout
    # the radix schedules of the BASELINE.json lengths are what the kernels were tuned and measured with:
    # adding radices (5, 7, 10, 14 serve other lengths) must not move them
    import re
    plans = {}
    for m in re.finditer(r"^  ([xz]) n= *(\d+) .*?plan((?: \d+)+)  ", r.stdout, re.M):
        plans[(m.group(1), int(m.group(2)))] = [int(v) for v in m.group(3).split()]
    want = {("x", 128): [16, 8], ("x", 384): [8, 8, 6], ("x", 1024): [16, 8, 8], ("x", 2048): [16, 16, 8],
            ("z", 128): [8, 4, 4], ("z", 384): [16, 6, 4], ("z", 768): [16, 12, 4],
            ("x", 640): [10, 8, 8], ("z", 640): [16, 10, 4], ("x", 896): [14, 8, 8], ("z", 448): [16, 7, 4]}
    for k, v in want.items():
        assert plans.get(k) == v, (k, plans.get(k), v)
