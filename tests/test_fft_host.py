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
