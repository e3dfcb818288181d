"""SURVEY 8(f) N1 -- the input pipeline in front of the step: a GPU-resident uint8 dataset, the reference's
DistributedSampler indices, gather and ToTensor's /255 fused into conv1's load.  Must be the same arithmetic, bit
for bit, as feeding the float batch the reference's DataLoader would have produced."""
import numpy as np
import pytest
import torch
from torch.utils.data.distributed import DistributedSampler

from oracle import convnet_oracle as O
from tests.refmodel import RefConvNet

pytestmark = pytest.mark.gpu


def _state():
    torch.manual_seed(3407)
    st = O.OracleState.from_state_dict(RefConvNet(None, "cpu").state_dict())
    bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]]).numpy()
    return st.flat_params().numpy(), bn


@pytest.mark.parametrize("dtype,B", [("fp16", 32), ("fp32", 32), ("fp16", 96), ("bf16", 96)])
def test_indexed_uint8_step_equals_float_batch_step(dtype, B):
    """B = 96 takes the large-batch kernels (conv1 and its weight gradient on tcgen05 read the uint8 rows too)."""
    from ddp_practice_b200 import Engine
    g = torch.Generator().manual_seed(5)
    n = 1000  # not a multiple of 32: the last batch is ragged, like 7500 % 32 in the reference at W=8
    images = torch.randint(0, 256, (n, 1, 28, 28), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (n,), generator=g)
    sampler = DistributedSampler(range(n), num_replicas=2, rank=1)  # ddp_main.py:130-132
    sampler.set_epoch(1)                                            # ddp_main.py:160
    order = torch.tensor(list(iter(sampler)), dtype=torch.int64)
    assert len(order) == 500
    params, bn = _state()
    a = Engine(dtype=dtype, batch_cap=B, use_scaler=dtype != "fp32")
    b = Engine(dtype=dtype, batch_cap=B, use_scaler=dtype != "fp32")
    for e in (a, b):
        e.load_state(params, bn)
    img_d, lab_d, ord_d = images.cuda(), labels.cuda(), order.cuda()
    for lo in range(0, 500, B):
        idx = ord_d[lo:lo + B].contiguous()
        a.train_step_indexed(img_d, lab_d, idx)
        # what the reference's pipeline hands to train(): ToTensor (uint8 -> float / 255), batch, .cuda()
        xb = (images[order[lo:lo + B]].float() / 255).cuda().contiguous()
        yb = labels[order[lo:lo + B]].cuda().contiguous()
        b.train_step(xb, yb)
        torch.cuda.synchronize()
        assert a.loss.item() == b.loss.item()
    assert torch.equal(a.params, b.params) and torch.equal(a.bn_buffers, b.bn_buffers)
    assert torch.equal(a.grads.view(torch.int32), b.grads.view(torch.int32))
    # test(): counts over the same shard, ddp_main.py:101-107
    a.reset_counts(); b.reset_counts()
    for lo in range(0, 500, B):
        idx = ord_d[lo:lo + B].contiguous()
        la = a.eval_batch_indexed(img_d, lab_d, idx).clone()
        xb = (images[order[lo:lo + B]].float() / 255).cuda().contiguous()
        lb = b.eval_batch(xb, labels[order[lo:lo + B]].cuda().contiguous())
        assert torch.equal(la, lb)
    assert a.read_counts() == b.read_counts() and a.read_counts()[1] == 500.0
    a.close(); b.close()
