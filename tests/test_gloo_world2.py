"""N > 1 host-side semantics on CPU: two real processes over gloo.
Pins what the multi-GPU path must reproduce -- sharding, the Reducer's mean, the rank-0 buffer
broadcast, the two-scalar reduce of test() -- with the oracle checked against a live 2-process
DistributedDataParallel run (not only the committed vectors)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn
from torch.utils.data.distributed import DistributedSampler

from oracle import convnet_oracle as O
from tests.refmodel import RefConvNet, RefTrainer, bn_buffers, flat_params


def test_distributed_sampler_known_answers():
    """the sharding every rank's data (and its test() counts) depends on."""
    s = DistributedSampler(range(10000), num_replicas=8, rank=3)
    idx = list(iter(s))
    assert len(idx) == 1250 and idx[:5] == [1917, 4869, 1426, 2751, 8950]
    s = DistributedSampler(range(60000), num_replicas=8, rank=3)
    s.set_epoch(1)
    idx = list(iter(s))
    assert len(idx) == 7500 and idx[:4] == [23645, 33583, 29135, 10586]
    # the ranks partition the permutation: rank r takes perm[r::W]
    parts = [list(iter(DistributedSampler(range(64), num_replicas=2, rank=r))) for r in range(2)]
    assert sorted(parts[0] + parts[1]) == list(range(64))


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    torch.manual_seed(50 + rank)
    model = nn.parallel.DistributedDataParallel(RefConvNet(None, "cpu"))
    init_p, init_bn = flat_params(model), bn_buffers(model)
    tr = RefTrainer(model, amp=False, device="cpu")
    B, steps = 8, 2
    x, y = O.synthetic_mnist(B * steps * world, seed=21)
    recs = []
    for s in range(steps):
        lo = (s * world + rank) * B
        recs.append(tr.step(x[lo:lo + B], y[lo:lo + B]))
    # test(): per-rank fp32 counters, reduce(SUM) to rank 0 (ddp_main.py:99-109)
    model.eval()
    xt, yt = O.synthetic_mnist(40 * world, seed=22)
    size, correct = torch.tensor(0.0), torch.tensor(0.0)
    for lo in range(0, 40, 16):
        images, labels = xt[rank::world][lo:lo + 16], yt[rank::world][lo:lo + 16]
        with torch.no_grad():
            outputs = model(images)
            size += images.size(0)
        correct += (outputs.argmax(1) == labels).type(torch.float).sum()
    dist.reduce(size, 0, op=dist.ReduceOp.SUM)
    dist.reduce(correct, 0, op=dist.ReduceOp.SUM)
    ret[rank] = {"init_p": init_p, "init_bn": init_bn, "recs": recs, "final_p": flat_params(model),
                 "final_bn": bn_buffers(model), "size": size.item(), "correct": correct.item()}
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_oracle_matches_live_two_process_gloo_ddp():
    world, port = 2, 29661
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
    r0, r1 = ret[0], ret[1]
    assert torch.equal(r0["init_p"], r1["init_p"]) and torch.equal(r0["init_bn"], r1["init_bn"])  # rank-0 broadcast
    st = O.OracleState(params={}, running_mean=[r0["init_bn"][0:16].clone(), r0["init_bn"][32:64].clone()],
                       running_var=[r0["init_bn"][16:32].clone(), r0["init_bn"][64:96].clone()])
    off = 0
    for n, shp in O.PARAM_SPECS:
        k = int(torch.tensor(shp).prod())
        st.params[n] = r0["init_p"][off:off + k].view(shp).clone()
        off += k
    B = 8
    x, y = O.synthetic_mnist(B * 2 * world, seed=21)
    for s in range(2):
        xs = [x[(s * world + r) * B:(s * world + r + 1) * B] for r in range(world)]
        ys = [y[(s * world + r) * B:(s * world + r + 1) * B] for r in range(world)]
        res = O.train_step(st, xs, ys, dtype="fp32", use_scaler=False, sync_bn=False)
        for r in range(world):
            assert res.losses[r] == pytest.approx(ret[r]["recs"][s]["loss"], rel=1e-5)
        assert torch.equal(r0["recs"][s]["grads"], r1["recs"][s]["grads"])  # one all-reduced bucket
        assert torch.allclose(res.grads, r0["recs"][s]["grads"], rtol=1e-4, atol=1e-6)
    assert torch.allclose(st.flat_params(), r0["final_p"], rtol=1e-5, atol=1e-7)
    assert torch.equal(r0["final_p"], r1["final_p"])
    # after the eval forward DDP has re-broadcast rank 0's BN buffers
    assert torch.equal(r0["final_bn"], r1["final_bn"])
    want_bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
    assert torch.allclose(want_bn, r0["final_bn"], rtol=1e-5, atol=1e-7)
    # test() counts: oracle's per-rank counting + rank-order sum == dist.reduce on rank 0
    xt, yt = O.synthetic_mnist(40 * world, seed=22)
    shards = [[(xt[r::world][lo:lo + 16], yt[r::world][lo:lo + 16]) for lo in range(0, 40, 16)] for r in range(world)]
    c, s = O.test_counts(st, shards, "fp32")
    assert (c, s) == (r0["correct"], r0["size"]) and s == 80.0


def _hook_worker(rank, world, port, ret):
    """The comm-hook contract (distributed.py:1987-2067) with the hook's reducer swapped for gloo: what must hold for
    ddp_practice_b200.comm_hook on any model -- >= 2 buckets, one Future per bucket, mean over ranks, replicas equal."""
    from ddp_practice_b200.comm_hook import B200AllreduceState, b200_allreduce_hook
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)

    def make():
        torch.manual_seed(7)
        return nn.Sequential(nn.Linear(64, 600), nn.ReLU(), nn.Linear(600, 600), nn.ReLU(), nn.Linear(600, 10))

    def gloo_mean(t):
        dist.all_reduce(t)
        t.div_(world)

    plain = nn.parallel.DistributedDataParallel(make(), bucket_cap_mb=0.2)
    hooked = nn.parallel.DistributedDataParallel(make(), bucket_cap_mb=0.2)
    state = B200AllreduceState(gloo_mean)
    hooked.register_comm_hook(state, b200_allreduce_hook)
    torch.manual_seed(100 + rank)
    x, y = torch.randn(16, 64), torch.randint(0, 10, (16,))
    # iteration 0 always runs with ONE bucket (distributed.py:1224-1225); the Reducer rebuilds its buckets under
    # bucket_cap_mb after it (reducer.hpp:299-301), so the second iteration is the one with several
    for it in range(2):
        for m in (plain, hooked):
            m.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
    same = all(torch.allclose(a.grad, b.grad, rtol=1e-6, atol=1e-8) for a, b in zip(plain.parameters(), hooked.parameters()))
    ret[rank] = {"same": same, "buckets": state.buckets_seen,
                 "grad0": torch.cat([p.grad.reshape(-1) for p in hooked.parameters()])}
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_comm_hook_contract_two_process_gloo():
    world, port = 2, 29671
    ret = mp.Manager().dict()
    mp.spawn(_hook_worker, args=(world, port, ret), nprocs=world, join=True)
    assert ret[0]["same"] and ret[1]["same"]
    assert ret[0]["buckets"] >= 3 and ret[0]["buckets"] == ret[1]["buckets"]  # 1 (iteration 0) + >= 2 (bucket_cap_mb)
    assert torch.equal(ret[0]["grad0"], ret[1]["grad0"])
