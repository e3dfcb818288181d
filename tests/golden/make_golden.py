"""Generate the golden vectors under tests/golden/ by running the REFERENCE code.

Run in the build container only (it imports /root/reference, which does not exist
on the GPU box):   python tests/golden/make_golden.py

The reference ships no tests and no golden files, so parity is
pinned on outputs of the reference itself:

  origin_fp32.npz   origin_main.ConvNet + the loop body of origin_main.py:60-66 on CPU
                    (`.cuda()` is the identity here), 3 steps, batch 8.
  amp_fp16.npz /    ddp_main.ConvNet + the loop body of ddp_main.py:86-93, with
  amp_bf16.npz      `torch.cuda.amp.autocast()` redirected to CPU autocast of the
                    same dtype and `torch.amp.GradScaler("cpu")`, 4 steps, batch 8.
  global_w4.npz     origin_main.ConvNet on a global batch of 32 = what DDP(W=4)+SyncBN
                    computes on 4 shards of 8 (SURVEY.md finding 6).
  ddp_gloo_w2.npz   ddp_main.ConvNet (plain BN, fp32) under real 2-process gloo
                    DistributedDataParallel: averaged grads, params, rank-0 buffers.
  test_counts.npz   the body of origin_main.test() on 96 images in 3 batches.
  origin_fp32_nesterov.npz / amp_fp16_momentum.npz
                    the same two loops with SGD(momentum, weight_decay, dampening / nesterov)
                    (`python tests/golden/make_golden.py sgd` writes only these).
  amp_fp16_accum.npz
                    ddp_main.ConvNet, fp16 autocast + GradScaler, TWO backward passes per optimizer step (gradient
                    accumulation, what DDP.no_sync() does per rank); `... make_golden.py accum` writes only this.
"""
import contextlib
import importlib.util
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle.convnet_oracle import PARAM_SPECS, synthetic_mnist  # noqa: E402  (names/shapes + data only)


def load_ref(name):
    spec = importlib.util.spec_from_file_location(f"ref_{name}", os.path.join(REF, f"{name}.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def sd_arrays(model, prefix):
    sd = {k.removeprefix("module."): v for k, v in model.state_dict().items()}
    out = {f"{prefix}params": torch.cat([sd[n].reshape(-1) for n, _ in PARAM_SPECS]).numpy().copy()}
    for li, l in enumerate(("layer1.1", "layer2.1")):
        out[f"{prefix}rm{li}"] = sd[f"{l}.running_mean"].numpy().copy()
        out[f"{prefix}rv{li}"] = sd[f"{l}.running_var"].numpy().copy()
        out[f"{prefix}nbt{li}"] = np.int64(sd[f"{l}.num_batches_tracked"])
    return out


def flat_grads(model):
    named = dict(model.named_parameters())
    named = {k.removeprefix("module."): v for k, v in named.items()}
    return torch.cat([named[n].grad.reshape(-1) for n, _ in PARAM_SPECS]).numpy().copy()


def sgd_arrays(kw):
    return {"sgd_" + k: np.float64(kw.get(k, d)) for k, d in
            (("lr", 1e-4), ("momentum", 0.0), ("weight_decay", 0.0), ("dampening", 0.0), ("nesterov", 0.0))}


def gen_origin(path, batch, steps, seed=3407, **sgd_kw):
    ref = load_ref("origin_main")
    torch.manual_seed(seed)
    model = ref.ConvNet()
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.SGD(model.parameters(), **{"lr": 1e-4, **sgd_kw})
    x, y = synthetic_mnist(batch * steps)
    out = sd_arrays(model, "init_")
    out["x"], out["y"] = x.numpy(), y.numpy()
    losses, grads = [], []
    model.train()
    for s in range(steps):
        images, labels = x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch]
        outputs = model(images)
        loss = criterion(outputs, labels)
        optimizer.zero_grad()
        loss.backward()
        grads.append(flat_grads(model))
        optimizer.step()
        losses.append(loss.item())
    out["losses"] = np.array(losses, dtype=np.float64)
    out["grads"] = np.stack(grads)
    out.update(sd_arrays(model, "final_"))
    out.update(sgd_arrays(sgd_kw))
    out["batch"], out["steps"] = np.int64(batch), np.int64(steps)
    np.savez_compressed(path, **out)


def gen_amp(path, dtype, batch=8, steps=4, seed=3407, **sgd_kw):
    ref = load_ref("ddp_main")
    # the reference's forward opens `torch.cuda.amp.autocast()`; on a CPU-only host that
    # context disables itself, so point it at CPU autocast of the dtype under test
    torch.cuda.amp.autocast = lambda: torch.autocast("cpu", dtype=dtype)
    torch.manual_seed(seed)
    model = ref.ConvNet()
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.SGD(model.parameters(), **{"lr": 1e-4, **sgd_kw})
    scaler = torch.amp.GradScaler("cpu")
    x, y = synthetic_mnist(batch * steps)
    out = sd_arrays(model, "init_")
    out["x"], out["y"] = x.numpy(), y.numpy()
    losses, scales, trackers, founds, grads = [], [], [], [], []
    model.train()
    for s in range(steps):
        images, labels = x[s * batch:(s + 1) * batch], y[s * batch:(s + 1) * batch]
        outputs = model(images)
        loss = criterion(outputs, labels)
        assert outputs.dtype == dtype and loss.dtype == dtype
        optimizer.zero_grad()
        scaler.scale(loss).backward()
        scaler.unscale_(optimizer)  # what scaler.step() does first; exposes tap T-c
        g = flat_grads(model)
        founds.append(0.0 if np.isfinite(g).all() else 1.0)
        grads.append(np.nan_to_num(g, nan=0.0, posinf=0.0, neginf=0.0))
        scaler.step(optimizer)
        scaler.update()
        losses.append(loss.item())
        scales.append(scaler.get_scale())
        trackers.append(int(scaler._growth_tracker.item()))
    out["losses"] = np.array(losses, dtype=np.float64)
    out["scales"] = np.array(scales, dtype=np.float64)
    out["trackers"] = np.array(trackers, dtype=np.int64)
    out["found_inf"] = np.array(founds, dtype=np.float64)
    out["grads"] = np.stack(grads)
    out.update(sd_arrays(model, "final_"))
    out.update(sgd_arrays(sgd_kw))
    out["batch"], out["steps"] = np.int64(batch), np.int64(steps)
    np.savez_compressed(path, **out)


def _gloo_worker(rank, world, port, batch, steps, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ref = load_ref("ddp_main")
    torch.cuda.amp.autocast = contextlib.nullcontext  # fp32 on CPU
    torch.manual_seed(1000 + rank)  # replicas start different; DDP broadcasts rank 0 (ddp_main.py:121)
    model = nn.parallel.DistributedDataParallel(ref.ConvNet())
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    x, y = synthetic_mnist(batch * steps * world)
    res = sd_arrays(model, "init_") if rank == 0 else {}
    grads, losses = [], []
    model.train()
    for s in range(steps):
        lo = (s * world + rank) * batch
        outputs = model(x[lo:lo + batch])
        loss = criterion(outputs, y[lo:lo + batch])
        optimizer.zero_grad()
        loss.backward()
        grads.append(flat_grads(model))
        optimizer.step()
        losses.append(loss.item())
    res[f"losses_r{rank}"] = np.array(losses)
    res[f"grads_r{rank}"] = np.stack(grads)
    # one eval forward triggers DDP's buffer broadcast from rank 0 (distributed.py:1590)
    model.eval()
    with torch.no_grad():
        model(x[:2])
    res.update(sd_arrays(model, f"final_r{rank}_"))
    ret[rank] = res
    dist.destroy_process_group()


def gen_gloo(path, batch=8, steps=2, world=2):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_gloo_worker, args=(world, 29611, batch, steps, ret), nprocs=world, join=True)
    out = {}
    for r in range(world):
        out.update(ret[r])
    x, y = synthetic_mnist(batch * steps * world)
    out["x"], out["y"] = x.numpy(), y.numpy()
    out["batch"], out["steps"], out["world"] = np.int64(batch), np.int64(steps), np.int64(world)
    np.savez_compressed(path, **out)


def gen_test_counts(path, seed=3407):
    ref = load_ref("origin_main")
    torch.manual_seed(seed)
    model = ref.ConvNet()
    x, y = synthetic_mnist(96 + 64, seed=11)
    model.train()
    with torch.no_grad():
        model(x[96:])  # move the running statistics off their init values
    out = sd_arrays(model, "init_")
    # body of origin_main.test() (origin_main.py:69-81)
    model.eval()
    size = torch.tensor(0.0)
    correct = torch.tensor(0.0)
    logits = []
    for lo, hi in ((0, 32), (32, 64), (64, 96)):
        images, labels = x[lo:hi], y[lo:hi]
        with torch.no_grad():
            outputs = model(images)
            size += images.size(0)
        correct += (outputs.argmax(1) == labels).type(torch.float).sum()
        logits.append(outputs)
    out["x"], out["y"] = x[:96].numpy(), y[:96].numpy()
    out["logits"] = torch.cat(logits).numpy()
    out["correct"], out["size"] = np.float32(correct.item()), np.float32(size.item())
    np.savez_compressed(path, **out)


def gen_amp_accum(path, dtype=torch.float16, batch=8, steps=4, seed=3407):
    """Gradient accumulation: the reference's model and loop body with TWO backward passes per optimizer
    step and no zero_grad() in between -- what `with model.no_sync():` around the first one does on every rank."""
    ref = load_ref("ddp_main")
    torch.cuda.amp.autocast = lambda: torch.autocast("cpu", dtype=dtype)
    torch.manual_seed(seed)
    model = ref.ConvNet()
    criterion = nn.CrossEntropyLoss()
    optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    scaler = torch.amp.GradScaler("cpu")
    x, y = synthetic_mnist(batch * steps * 2)
    out = sd_arrays(model, "init_")
    out["x"], out["y"] = x.numpy(), y.numpy()
    losses, scales, trackers, founds, grads = [], [], [], [], []
    model.train()
    for s in range(steps):
        optimizer.zero_grad()
        for m in range(2):
            lo = (2 * s + m) * batch
            loss = criterion(model(x[lo:lo + batch]), y[lo:lo + batch])
            scaler.scale(loss).backward()
            losses.append(loss.item())
        scaler.unscale_(optimizer)
        g = flat_grads(model)
        founds.append(0.0 if np.isfinite(g).all() else 1.0)
        grads.append(np.nan_to_num(g, nan=0.0, posinf=0.0, neginf=0.0))
        scaler.step(optimizer)
        scaler.update()
        scales.append(scaler.get_scale())
        trackers.append(int(scaler._growth_tracker.item()))
    out["losses"] = np.array(losses, dtype=np.float64)
    out["scales"] = np.array(scales, dtype=np.float64)
    out["trackers"] = np.array(trackers, dtype=np.int64)
    out["found_inf"] = np.array(founds, dtype=np.float64)
    out["grads"] = np.stack(grads)
    out.update(sd_arrays(model, "final_"))
    out["batch"], out["steps"] = np.int64(batch), np.int64(steps)
    np.savez_compressed(path, **out)


def gen_sgd_knobs():
    """the reference's model / loop body with the other torch.optim.SGD knobs switched on (the
    reference itself only ever passes lr; the optimizer class is the same)."""
    gen_origin(os.path.join(HERE, "origin_fp32_nesterov.npz"), batch=8, steps=4, lr=1e-2, momentum=0.9,
               weight_decay=5e-4, nesterov=True)
    gen_amp(os.path.join(HERE, "amp_fp16_momentum.npz"), torch.float16, steps=5, lr=1e-3, momentum=0.9,
            dampening=0.1, weight_decay=1e-3)


if __name__ == "__main__":
    torch.set_num_threads(1)
    torch.use_deterministic_algorithms(True)
    if sys.argv[1:] == ["sgd"]:  # only the N3 vectors (leaves the others byte-identical)
        gen_sgd_knobs()
        raise SystemExit(0)
    if sys.argv[1:] == ["accum"]:
        gen_amp_accum(os.path.join(HERE, "amp_fp16_accum.npz"))
        raise SystemExit(0)
    gen_origin(os.path.join(HERE, "origin_fp32.npz"), batch=8, steps=3)
    gen_origin(os.path.join(HERE, "global_w4.npz"), batch=32, steps=2)
    gen_amp(os.path.join(HERE, "amp_fp16.npz"), torch.float16)
    gen_amp(os.path.join(HERE, "amp_bf16.npz"), torch.bfloat16)
    gen_test_counts(os.path.join(HERE, "test_counts.npz"))
    gen_gloo(os.path.join(HERE, "ddp_gloo_w2.npz"))
    gen_sgd_knobs()
    gen_amp_accum(os.path.join(HERE, "amp_fp16_accum.npz"))
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
