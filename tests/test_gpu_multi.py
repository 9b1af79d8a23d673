"""Multi-GPU parity under pytest: tests/mgpu_check.py (every rank steps its kx / y-plane slab of the golden cases and
compares it with the dumps of the reference build; operator-level convolution included) launched with torchrun on the
GPUs of the box, with the peer-copy exchange (bulk NVLink copies into CUDA-IPC-mapped buffers) and with the NCCL send/recv
exchange.  Skipped on a one-GPU box; bench.py carries the same check into every N>1 line (`parity_vs_golden`)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("peer", ["1", "0"])
def test_slab_parity_on_all_gpus(peer):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    n = min(n, 4)
    env = dict(os.environ, CHAN_PEER=peer)
    port = 29540 + (1 if peer == "1" else 0)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu_check.py")], capture_output=True, text=True, timeout=900, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("multi-GPU parity OK") == n, r.stdout[-2000:]
