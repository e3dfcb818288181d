"""torchrun worker: the fused step on W GPUs against the reference's own DDP + SyncBN + AMP stack
(torch.nn.parallel.DistributedDataParallel over NCCL) running in the same processes.

Launched by tests/test_gpu_multi.py:  torchrun --nproc-per-node W ddp_parity_worker.py <dtype>
Every rank checks its own numbers; any assertion failure exits non-zero."""
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from ddp_practice_b200 import Engine  # noqa: E402
from oracle import convnet_oracle as O  # noqa: E402
from tests.helpers import assert_grads_close, grad_errors  # noqa: E402
from tests.refmodel import RefConvNet, RefTrainer, bn_buffers, flat_params, nbt  # noqa: E402

TDT = {"fp32": None, "fp16": torch.float16, "bf16": torch.bfloat16}


def main():
    dtype = sys.argv[1] if len(sys.argv) > 1 else "fp16"
    sync_bn = (sys.argv[2] != "plainbn") if len(sys.argv) > 2 else True
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    amp = dtype != "fp32"
    B = int(sys.argv[3]) if len(sys.argv) > 3 else 32
    # the reference's construction sequence, ddp_main.py:117-126 (replicas start different: DDP broadcasts rank 0)
    torch.manual_seed(1000 + rank)
    model = RefConvNet(TDT[dtype], "cuda").cuda()
    if sync_bn:
        model = nn.SyncBatchNorm.convert_sync_batchnorm(model)
    model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
    tr = RefTrainer(model, amp)

    eng = Engine(dtype=dtype, batch_cap=B, rank=rank, world=world, device=local, use_scaler=amp, sync_bn=sync_bn)
    eng.connect()
    eng.load_state(flat_params(model).numpy(), bn_buffers(model).numpy())

    # DistributedSampler-style shard: rank r takes samples r::world of a fixed permutation
    steps, tail = 5, 12
    n_total = (steps * B + tail) * world
    x, y = O.synthetic_mnist(n_total)
    x, y = x[rank::world].cuda(), y[rank::world].cuda()
    bounds = [(i * B, (i + 1) * B) for i in range(steps)] + [(steps * B, steps * B + tail)]
    tol = 2e-2 if amp else 1e-3
    worst = 0.0
    for s, (lo, hi) in enumerate(bounds):
        xs, ys = x[lo:hi].contiguous(), y[lo:hi].contiguous()
        ref = tr.step(xs, ys)
        eng.train_step(xs, ys)
        torch.cuda.synchronize()
        eng.check_device()
        assert eng.found_inf.item() == ref["found_inf"], (rank, s)
        if amp:
            assert eng.scale.item() == ref["scale"] and eng.growth_tracker.item() == ref["tracker"], (rank, s)
        assert abs(eng.loss.item() - ref["loss"]) <= tol * abs(ref["loss"]), (rank, s, eng.loss.item(), ref["loss"])
        if ref["found_inf"] == 0.0:
            errs = grad_errors(eng.grads.cpu(), ref["grads"])
            worst = max(worst, max(errs.values()))
            # routing flips (tests/helpers.py) are not counted here: bound for steps that may contain them
            assert_grads_close(eng.grads.cpu(), ref["grads"], 6e-2 if amp else 2e-2, f"rank {rank} step {s}")
        # every replica must hold bit-identical parameters and gradients
        mine = torch.cat([eng.params, eng.grads, eng.bn_buffers]).clone().view(torch.int32)  # bit patterns (NaN-safe)
        other = mine.clone()
        dist.broadcast(other, 0)
        assert torch.equal(mine, other), f"rank {rank} diverged from rank 0 at step {s}"
    got = eng.read_state()
    assert torch.allclose(torch.from_numpy(got["params"]), flat_params(model), rtol=1e-4, atol=2e-6)
    # the reference's buffers are compared after its next forward-time broadcast (plain BN) / directly (SyncBN)
    model.eval()
    with torch.no_grad():
        model(x[:2])
    assert torch.allclose(torch.from_numpy(got["bn_buffers"]), bn_buffers(model), rtol=2e-3 if amp else 1e-5, atol=1e-6)
    assert list(got["nbt"]) == nbt(model)

    # ---- test(): ddp_main.py:96-112 on both stacks
    xt, yt = O.synthetic_mnist(1000 * world, seed=9)
    xt, yt = xt[rank::world].cuda(), yt[rank::world].cuda()
    size = torch.tensor(0.0).cuda()
    correct = torch.tensor(0.0).cuda()
    mine_c = torch.tensor(0.0).cuda()
    eng.reset_counts()
    for lo in range(0, 1000, B):
        images, labels = xt[lo:lo + B].contiguous(), yt[lo:lo + B].contiguous()
        with torch.no_grad():
            outputs = model(images)
            size += images.size(0)
        correct += (outputs.argmax(1) == labels).type(torch.float).sum()
        lg = eng.eval_batch(images, labels)
        mine_c += (lg.argmax(1) == labels).type(torch.float).sum()
    dist.reduce(size, 0, op=dist.ReduceOp.SUM)
    dist.reduce(correct, 0, op=dist.ReduceOp.SUM)
    dist.reduce(mine_c, 0, op=dist.ReduceOp.SUM)
    eng.reduce_counts()
    c, sz = eng.read_counts()
    eng.check_device()
    if rank == 0:
        assert sz == size.item() == 1000.0 * world
        assert c == mine_c.item()                 # P2P reduce == dist.reduce of the same per-rank counts
        assert abs(c - correct.item()) <= 3 * world  # vs the reference's own predictions (near-tie flips only)
        print(json.dumps({"ok": True, "dtype": dtype, "world": world, "sync_bn": sync_bn, "worst_grad_rel_l2": worst,
                          "correct": c, "size": sz, "ref_correct": correct.item()}), flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
