"""Multi-GPU parity (needs >= 2 GPUs on one node; skipped otherwise): the fused step with its
NVLink peer-memory exchanges against torch DDP + SyncBN + AMP over NCCL, one process per GPU."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "mgpu", "ddp_parity_worker.py")


def _run(world, dtype, mode="syncbn", port=29731, batch=32):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(port), WORKER, dtype, mode, str(batch)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    line = [l for l in res.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


def _need(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} GPUs")


@pytest.mark.parametrize("dtype", ["fp16", "bf16", "fp32"])
def test_two_gpu_ddp_syncbn_amp_parity(dtype):
    _need(2)
    out = _run(2, dtype)
    assert out["ok"] and out["size"] == 2000.0


@pytest.mark.parametrize("batch", [48, 80])
def test_two_gpu_other_batch_sizes(batch):
    """48 per rank: BN1 statistics through tagged rows, BN2 through plain deferred rows (96 partial rows); 80 per rank:
    the large-batch kernels (tcgen05 conv1, statistics finalised and pushed by the producer's last CTA)."""
    _need(2)
    assert _run(2, "fp16", batch=batch)["ok"]


def test_two_gpu_plain_bn_ddp_parity():
    _need(2)
    assert _run(2, "fp16", "plainbn")["ok"]


def test_all_gpus_ddp_syncbn_amp_parity():
    n = torch.cuda.device_count()
    if n < 4:
        pytest.skip("needs >= 4 GPUs")
    out = _run(n, "fp16")
    assert out["ok"] and out["size"] == 1000.0 * n
