"""torchrun worker: the fused step on W GPUs against the reference's own DDP + SyncBN + AMP stack
(torch.nn.parallel.DistributedDataParallel over NCCL) running in the same processes.

Launched by tests/test_gpu_multi.py:
    torchrun --nproc-per-node W ddp_parity_worker.py <dtype> <syncbn|plainbn> <batch> <test images per rank>
Every rank checks its own numbers; any assertion failure exits non-zero.  Rank 0 prints one JSON line with the
per-step evidence (routing flips, worst relative L2 gradient error, tolerance applied) and the test() accounting."""
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
from tests.helpers import assert_grads_close, count_flips, grad_errors, routing_of  # noqa: E402
from tests.refmodel import RefConvNet, RefTrainer, bn_buffers, flat_params, nbt  # noqa: E402

TDT = {"fp32": None, "fp16": torch.float16, "bf16": torch.bfloat16}
AMP_TOL, FP32_TOL = 2e-2, 1e-3          # north_star's 2e-2 under autocast; fp32 vs cuDNN's own fp32 algorithms: 1e-3
AMP_FLIP_TOL, FP32_FLIP_TOL = 6e-2, 2e-2  # steps where the two stacks disagree on >= 1 ReLU / pool routing decision
MAX_FLIP_FRACTION = 1e-3                # of all pooling windows of the global batch


def global_bn_output(y_local, gamma, beta, dtype, sync_bn, world):
    """BN(train) output of this rank's conv output `y_local` under the statistics the step used: this rank's own
    (plain BN) or the Chan merge of all ranks' in rank order (SyncBN), with the oracle's formulas."""
    mean, invstd, n = O._local_stats(y_local)
    if sync_bn and world > 1:
        pack = torch.cat([mean, invstd]).cuda()
        allp = [torch.empty_like(pack) for _ in range(world)]
        dist.all_gather(allp, pack)
        C = mean.numel()
        mean, invstd, _, _ = O.merge_stats([p[:C].cpu() for p in allp], [p[C:].cpu() for p in allp], [n] * world)
    return O._bn_apply(y_local, mean, invstd, gamma, beta, O._DT[dtype])


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
    n_test = int(sys.argv[4]) if len(sys.argv) > 4 else 1000
    # the reference's construction sequence, ddp_main.py:117-126 (replicas start different: DDP broadcasts rank 0)
    torch.manual_seed(1000 + rank)
    model = RefConvNet(TDT[dtype], "cuda").cuda()
    if sync_bn:
        model = nn.SyncBatchNorm.convert_sync_batchnorm(model)
    model = nn.parallel.DistributedDataParallel(model, device_ids=[local])
    bn_out = {}
    model.module.layer1[1].register_forward_hook(lambda m, i, o: bn_out.__setitem__(1, o.detach().float().cpu()))
    model.module.layer2[1].register_forward_hook(lambda m, i, o: bn_out.__setitem__(2, o.detach().float().cpu()))
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
    tol, flip_tol = (AMP_TOL, AMP_FLIP_TOL) if amp else (FP32_TOL, FP32_FLIP_TOL)
    per_step = []
    for s, (lo, hi) in enumerate(bounds):
        xs, ys = x[lo:hi].contiguous(), y[lo:hi].contiguous()
        sd = {k.removeprefix("module."): v.detach().float().cpu().clone() for k, v in model.state_dict().items()}
        ref = tr.step(xs, ys)
        eng.train_step(xs, ys)
        torch.cuda.synchronize()
        eng.check_device()
        assert eng.found_inf.item() == ref["found_inf"], (rank, s)
        if amp:
            assert eng.scale.item() == ref["scale"] and eng.growth_tracker.item() == ref["tracker"], (rank, s)
        assert abs(eng.loss.item() - ref["loss"]) <= tol * abs(ref["loss"]), (rank, s, eng.loss.item(), ref["loss"])
        # routing decisions (ReLU mask, max-pool winner) of the two stacks on THIS rank's images, under the statistics
        # of the global batch; summed over ranks because the all-reduced gradient carries every rank's flips
        n = hi - lo
        flips = (count_flips(routing_of(global_bn_output(eng.read_activation("y1", n), sd["layer1.1.weight"], sd["layer1.1.bias"],
                                                         dtype, sync_bn, world)), routing_of(bn_out[1])) +
                 count_flips(routing_of(global_bn_output(eng.read_activation("y2", n), sd["layer2.1.weight"], sd["layer2.1.bias"],
                                                         dtype, sync_bn, world)), routing_of(bn_out[2])))
        ft = torch.tensor([flips], device="cuda")
        dist.all_reduce(ft)
        flips = int(ft.item())
        assert flips <= MAX_FLIP_FRACTION * n * world * (16 * 196 + 32 * 49), (rank, s, flips)
        rec = {"step": s, "batch": n, "found_inf": ref["found_inf"], "flips": flips, "worst_rel_l2": None,
               "tol": None, "loss": eng.loss.item(), "ref_loss": ref["loss"]}
        if ref["found_inf"] == 0.0:
            errs = grad_errors(eng.grads.cpu(), ref["grads"])
            gtol = tol if flips == 0 else flip_tol  # the stated tolerance holds on every step without routing flips
            rec["worst_rel_l2"], rec["tol"] = max(errs.values()), gtol
            assert_grads_close(eng.grads.cpu(), ref["grads"], gtol, f"rank {rank} step {s} ({flips} routing flips)")
        per_step.append(rec)
        # every replica must hold bit-identical parameters and gradients
        mine = torch.cat([eng.params, eng.grads, eng.bn_buffers]).clone().view(torch.int32)  # bit patterns (NaN-safe)
        other = mine.clone()
        dist.broadcast(other, 0)
        assert torch.equal(mine, other), f"rank {rank} diverged from rank 0 at step {s}"
    # ---- DDP.no_sync() / gradient accumulation (distributed.py:1441-1470): one micro-step whose gradient
    # stays local, then a synchronising step that all-reduces the sum -- twice
    xa, ya = O.synthetic_mnist(4 * B * world, seed=21)
    xa, ya = xa[rank::world].cuda(), ya[rank::world].cuda()
    acc_steps = []
    for s in range(2):
        x1, y1 = xa[(2 * s) * B:(2 * s + 1) * B].contiguous(), ya[(2 * s) * B:(2 * s + 1) * B].contiguous()
        x2, y2 = xa[(2 * s + 1) * B:(2 * s + 2) * B].contiguous(), ya[(2 * s + 1) * B:(2 * s + 2) * B].contiguous()
        model.train()
        tr.optimizer.zero_grad()
        with model.no_sync():
            l1 = tr.criterion(model(x1), y1)
            (tr.scaler.scale(l1) if amp else l1).backward()
        l2 = tr.criterion(model(x2), y2)
        (tr.scaler.scale(l2) if amp else l2).backward()
        if amp:
            tr.scaler.unscale_(tr.optimizer)
        from tests.refmodel import flat_grads
        g_ref = flat_grads(model)
        finite = bool(torch.isfinite(g_ref).all())
        if amp:
            tr.scaler.step(tr.optimizer)
            tr.scaler.update()
        else:
            tr.optimizer.step()
        eng.train_step_accumulate(x1, y1)
        eng.train_step(x2, y2)
        torch.cuda.synchronize()
        eng.check_device()
        assert eng.found_inf.item() == (0.0 if finite else 1.0), (rank, "no_sync", s)
        if amp:
            assert eng.scale.item() == tr.scaler.get_scale() and eng.growth_tracker.item() == int(tr.scaler._growth_tracker.item())
        assert abs(eng.loss.item() - float(l2.item())) <= tol * abs(float(l2.item()))
        if finite:
            errs = grad_errors(eng.grads.cpu(), g_ref)
            acc_steps.append(max(errs.values()))
            assert_grads_close(eng.grads.cpu(), g_ref, flip_tol, f"rank {rank} no_sync step {s}")  # flips not counted here
        mine = torch.cat([eng.params, eng.grads, eng.bn_buffers]).clone().view(torch.int32)
        other = mine.clone()
        dist.broadcast(other, 0)
        assert torch.equal(mine, other), f"rank {rank} diverged from rank 0 in no_sync step {s}"

    got = eng.read_state()
    assert torch.allclose(torch.from_numpy(got["params"]), flat_params(model), rtol=1e-4, atol=2e-6)
    # the reference's buffers are compared after its next forward-time broadcast (plain BN) / directly (SyncBN)
    model.eval()
    with torch.no_grad():
        model(x[:2])
    assert torch.allclose(torch.from_numpy(got["bn_buffers"]), bn_buffers(model), rtol=2e-3 if amp else 1e-5, atol=1e-6)
    assert list(got["nbt"]) == nbt(model)

    # ---- test(): ddp_main.py:96-112 on both stacks.  What is bit-exact: `size`; the count of OUR logits (in-kernel
    # argmax + count == torch's argmax on the same logits) and the 2-scalar reduce to rank 0 (== dist.reduce).  Against
    # the reference's OWN predictions the counts can only differ where its top-2 logits are within 4 ulp of the
    # autocast dtype (cuBLASLt's and our Linear differ in the last bit): every disagreement must be such a near-tie,
    # and |correct - ref_correct| is bounded by their number.
    xt, yt = O.synthetic_mnist(n_test * world, seed=9)
    xt, yt = xt[rank::world].cuda(), yt[rank::world].cuda()
    size = torch.tensor(0.0).cuda()
    correct = torch.tensor(0.0).cuda()
    mine_c = torch.tensor(0.0).cuda()
    tie_flips = torch.tensor(0.0).cuda()
    hard_flips = torch.tensor(0.0).cuda()
    min_margin = torch.tensor(float("inf")).cuda()
    eps = torch.finfo(TDT[dtype] or torch.float32).eps
    eng.reset_counts()
    Be = min(B, 32)  # ddp_main.py:151 evaluates with the training batch size; the ragged last batch is included
    for lo in range(0, n_test, Be):
        images, labels = xt[lo:lo + Be].contiguous(), yt[lo:lo + Be].contiguous()
        with torch.no_grad():
            outputs = model(images)
            size += images.size(0)
        correct += (outputs.argmax(1) == labels).type(torch.float).sum()
        lg = eng.eval_batch(images, labels)
        mine_c += (lg.argmax(1) == labels).type(torch.float).sum()
        top2 = outputs.float().topk(2, dim=1).values
        margin = (top2[:, 0] - top2[:, 1]).abs()
        near_tie = margin <= 4 * eps * top2[:, 0].abs().clamp_min(1e-3)
        differ = lg.argmax(1) != outputs.argmax(1)
        tie_flips += (differ & near_tie).sum()
        hard_flips += (differ & ~near_tie).sum()
        min_margin = torch.minimum(min_margin, (margin / top2[:, 0].abs().clamp_min(1e-3)).min())
    dist.reduce(size, 0, op=dist.ReduceOp.SUM)
    dist.reduce(correct, 0, op=dist.ReduceOp.SUM)
    dist.reduce(mine_c, 0, op=dist.ReduceOp.SUM)
    dist.reduce(tie_flips, 0, op=dist.ReduceOp.SUM)
    dist.reduce(hard_flips, 0, op=dist.ReduceOp.SUM)
    dist.reduce(min_margin, 0, op=dist.ReduceOp.MIN)
    eng.reduce_counts()
    c, sz = eng.read_counts()
    eng.check_device()
    if rank == 0:
        assert sz == size.item() == float(n_test * world)
        assert c == mine_c.item()                 # P2P reduce == dist.reduce of the same per-rank counts (bit-exact)
        assert hard_flips.item() == 0.0           # a prediction may differ from the reference's only at a near-tie
        assert abs(c - correct.item()) <= tie_flips.item()
        print(json.dumps({"ok": True, "dtype": dtype, "world": world, "sync_bn": sync_bn, "batch": B,
                          "worst_grad_rel_l2": max([r["worst_rel_l2"] or 0.0 for r in per_step]),
                          "steps_needing_flip_tol": sum(1 for r in per_step if r["flips"] > 0 and r["tol"] is not None),
                          "per_step": per_step, "no_sync_worst_rel_l2": acc_steps, "correct": c, "size": sz, "ref_correct": correct.item(),
                          "near_tie_prediction_flips": tie_flips.item(), "non_tie_prediction_flips": hard_flips.item(),
                          "min_top2_rel_margin": min_margin.item()}), flush=True)
    eng.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
