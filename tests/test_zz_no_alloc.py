"""The allocation-free entry point, `ddpstep_conv1_bnstats_noalloc`, against the ordinary one.

Who needs it: a host that may not allocate device memory when it wants conv1 + BN statistics
-- every byte of the GPU already belongs to some other allocator (a caching pool, an arena
reserved at start-up), or it sits in a loop that cannot afford cudaMalloc's device-wide stall.
Such a caller
cannot create a context (that allocates); this entry point takes only caller-owned buffers
and keeps its scratch in a fixed array inside the library (limit: 1024 images per call).

The check: run the same batch through a normal context (`ddpstep_train_step`, whose conv1
output is read back with `ddpstep_read_activation`) and through the new entry point, and
compare -- twice in a row with different batches, because the second call finds the scratch
array as the first one left it.  Collected last on purpose (file name), so a run that stops
at the first failure has already run everything else."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import convnet_oracle as O
from tests.refmodel import RefConvNet

pytestmark = pytest.mark.gpu


def _call(lib, params, x, dtype_id, tdt, eps=1e-5):
    B = x.shape[0]
    y1 = torch.empty((B, 16, 28, 28), dtype=tdt, device="cuda")
    stats = torch.empty(32, dtype=torch.float32, device="cuda")
    rc = lib.ddpstep_conv1_bnstats_noalloc(params.data_ptr(), x.data_ptr(), B, dtype_id, 1, eps, y1.data_ptr(),
                                           stats.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, lib.ddpstep_last_error(None)
    torch.cuda.synchronize()
    return y1, stats


@pytest.mark.parametrize("dtype,dtype_id,tdt", [("fp16", 1, torch.float16), ("fp32", 0, torch.float32),
                                                ("bf16", 2, torch.bfloat16)])
def test_no_alloc_entry_point_matches_context_path_twice_in_a_row(dtype, dtype_id, tdt):
    from ddp_practice_b200 import Engine, _lib
    lib = _lib.load()
    torch.manual_seed(3407)
    st = O.OracleState.from_state_dict(RefConvNet(None, "cpu").state_dict())
    bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]]).numpy()
    x, y = O.synthetic_mnist(32 + 12, seed=77)
    params = st.flat_params().cuda()
    # two calls in a row, different batch sizes and data: 32 images, then a ragged 12
    for lo, hi in ((0, 32), (32, 44)):
        xs, ys = x[lo:hi].cuda().contiguous(), y[lo:hi].cuda().contiguous()
        n = hi - lo
        # the existing entry point: a context, fresh parameters, one step, then read conv1's output
        eng = Engine(dtype=dtype, batch_cap=32, use_scaler=dtype != "fp32")
        eng.load_state(st.flat_params().numpy(), bn)
        eng.train_step(xs, ys)
        torch.cuda.synchronize()
        y1_ref = eng.read_activation("y1", n)
        rm1 = eng.bn_buffers[:16].cpu()  # running_mean after one step from zero = 0.1 * batch mean
        eng.close()
        y1, stats = _call(lib, params, xs, dtype_id, tdt)
        assert torch.equal(y1.float().cpu(), y1_ref), f"conv1 output differs (batch {n})"
        mean, invstd = stats[:16].cpu(), stats[16:].cpu()
        assert torch.allclose(mean, rm1 * 10.0, rtol=1e-5, atol=1e-7)
        # and against statistics taken from the existing path's conv1 output
        m_ref, is_ref, _ = O._local_stats(y1_ref)
        assert torch.allclose(mean, m_ref, rtol=1e-5, atol=1e-7)
        assert torch.allclose(invstd, is_ref, rtol=1e-5)


@pytest.mark.parametrize("dtype,dtype_id,tdt", [("fp16", 1, torch.float16), ("bf16", 2, torch.bfloat16)])
def test_no_alloc_entry_point_large_batches_up_to_the_limit(dtype, dtype_id, tdt, monkeypatch):
    """Above 296 images this entry point launches one CTA per image (one partial row each, up to the 1024 rows of the
    fixed scratch) while a context grid-strides over 296 CTAs: 300 images, then the limit itself, 1024, in a row.
    The context is told to run conv1 with the same CUDA-core kernel (DDPSTEP_CONV1_TC=0; its default above 64 images is
    the tensor-core kernel), so the conv output must agree bit for bit; the statistics are sums in a different order
    and are compared at 1e-5."""
    from ddp_practice_b200 import Engine, _lib
    monkeypatch.setenv("DDPSTEP_CONV1_TC", "0")
    lib = _lib.load()
    torch.manual_seed(3407)
    st = O.OracleState.from_state_dict(RefConvNet(None, "cpu").state_dict())
    bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]]).numpy()
    x, y = O.synthetic_mnist(300 + 1024, seed=78)
    params = st.flat_params().cuda()
    eng = Engine(dtype=dtype, batch_cap=1024)
    for lo, hi in ((0, 300), (300, 1324)):
        xs, ys = x[lo:hi].cuda().contiguous(), y[lo:hi].cuda().contiguous()
        n = hi - lo
        eng.load_state(st.flat_params().numpy(), bn)
        eng.train_step(xs, ys)
        torch.cuda.synchronize()
        y1_ref = eng.read_activation("y1", n)
        y1, stats = _call(lib, params, xs, dtype_id, tdt)
        assert torch.equal(y1.float().cpu(), y1_ref), f"conv1 output differs (batch {n})"
        m_ref, is_ref, _ = O._local_stats(y1_ref)
        assert torch.allclose(stats[:16].cpu(), m_ref, rtol=1e-5, atol=1e-6)
        assert torch.allclose(stats[16:].cpu(), is_ref, rtol=1e-5)
    eng.close()


def test_no_alloc_entry_point_rejects_misaligned_buffers():
    """A caller carving buffers out of its own arena may hand over an odd offset: refused up front (the kernel's 16-byte
    accesses would otherwise raise a sticky fault in the caller's CUDA context)."""
    from ddp_practice_b200 import _lib
    lib = _lib.load()
    p = torch.zeros(416, device="cuda")
    x = torch.zeros(8 * 784 + 4, device="cuda")
    y1 = torch.zeros(8 * 16 * 784 + 4, device="cuda")
    out = torch.zeros(32, device="cuda")
    rc = lib.ddpstep_conv1_bnstats_noalloc(p.data_ptr(), x.data_ptr() + 4, 8, 0, 1, 1e-5, y1.data_ptr(), out.data_ptr(), None)
    assert rc == -1 and b"aligned" in lib.ddpstep_last_error(None)
    rc = lib.ddpstep_conv1_bnstats_noalloc(p.data_ptr(), x.data_ptr(), 8, 0, 1, 1e-5, y1.data_ptr() + 8, out.data_ptr(), None)
    assert rc == -1 and b"aligned" in lib.ddpstep_last_error(None)
    torch.cuda.synchronize()  # nothing was launched, nothing faulted


def test_no_alloc_entry_point_rejects_oversize_batch():
    from ddp_practice_b200 import _lib
    lib = _lib.load()
    p = torch.zeros(416, device="cuda")
    x = torch.zeros(8, 784, device="cuda")
    out = torch.zeros(32, device="cuda")
    rc = lib.ddpstep_conv1_bnstats_noalloc(p.data_ptr(), x.data_ptr(), 1025, 0, 1, 1e-5, x.data_ptr(), out.data_ptr(), None)
    assert rc == -1
    assert b"1024" in lib.ddpstep_last_error(None)
