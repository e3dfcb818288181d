"""The drop-in shim on a GPU: the reference's loop written verbatim against shim objects."""
import numpy as np
import pytest
import torch

from oracle import convnet_oracle as O
from tests.helpers import assert_grads_close
from tests.refmodel import RefConvNet, RefTrainer, bn_buffers, flat_grads, flat_params

pytestmark = pytest.mark.gpu


def _reference_style_loop(model, batches, criterion, optimizer, scaler):
    model.train()
    for images, labels in batches:  # ddp_main.py:85-93
        images = images.cuda()
        labels = labels.cuda()
        outputs = model(images)
        loss = criterion(outputs, labels)
        optimizer.zero_grad()
        scaler.scale(loss).backward()
        scaler.step(optimizer)
        scaler.update()
    return loss


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_shim_runs_the_reference_loop_and_matches_torch(dtype):
    from ddp_practice_b200 import Engine
    from ddp_practice_b200.shim import GradScaler, nn
    torch.manual_seed(3407)
    ref_model = RefConvNet(dtype, "cuda").cuda()
    torch.manual_seed(3407)
    model = RefConvNet(dtype, "cuda").cuda()
    init_p, init_bn = flat_params(model).numpy(), bn_buffers(model).numpy()
    # construction sequence of ddp_main.py:117-126
    model = nn.SyncBatchNorm.convert_sync_batchnorm(model)
    model = nn.parallel.DistributedDataParallel(model, device_ids=[0])
    criterion = nn.CrossEntropyLoss().cuda()
    optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    scaler = GradScaler()
    assert model.amp_dtype == {torch.float16: "fp16", torch.bfloat16: "bf16"}[dtype]
    x, y = O.synthetic_mnist(32 * 4 + 12)
    batches = [(x[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32]) for i in range(4)] + [(x[128:], y[128:])]
    loss = _reference_style_loop(model, batches, criterion, optimizer, scaler)
    torch.cuda.synchronize()

    # (1) identical bits to driving the C ABI directly
    e = Engine(dtype=model.amp_dtype, batch_cap=32)
    e.load_state(init_p, init_bn)
    for xb, yb in batches:
        e.train_step(xb.cuda().contiguous(), yb.cuda().contiguous())
    torch.cuda.synchronize()
    assert torch.equal(flat_params(model), e.params.cpu())
    assert torch.equal(bn_buffers(model), e.bn_buffers.cpu())
    assert loss.item() == e.loss.item()
    assert scaler.get_scale() == e.scale.item()

    # (2) tolerance-level agreement with the reference stack
    tr = RefTrainer(ref_model, True)
    for xb, yb in batches:
        rec = tr.step(xb.cuda(), yb.cuda())
    assert scaler.state_dict() == {"scale": rec["scale"], "growth_factor": 2.0, "backoff_factor": 0.5,
                                   "growth_interval": 2000, "_growth_tracker": rec["tracker"]}
    assert loss.item() == pytest.approx(rec["loss"], rel=2e-2)
    assert_grads_close(flat_grads(model), rec["grads"], 6e-2, "shim p.grad (tap T-c)")
    assert np.allclose(flat_params(model).numpy(), flat_params(ref_model).numpy(), rtol=1e-4, atol=2e-6)

    # (3) checkpoint surface, ddp_main.py:165-169
    sd = model.state_dict()
    assert list(sd.keys()) == ["module." + k for k in ref_model.state_dict().keys()] and len(sd) == 16
    assert int(sd["module.layer1.1.num_batches_tracked"]) == 5

    # (4) test(), ddp_main.py:96-107
    model.eval(); ref_model.eval()
    size = torch.tensor(0.0).cuda(); correct = torch.tensor(0.0).cuda(); ref_correct = torch.tensor(0.0).cuda()
    xt, yt = O.synthetic_mnist(200, seed=3)
    for lo in range(0, 200, 32):
        images, labels = xt[lo:lo + 32].cuda(), yt[lo:lo + 32].cuda()
        with torch.no_grad():
            outputs = model(images)
            size += images.size(0)
            ref_correct += (ref_model(images).argmax(1) == labels).type(torch.float).sum()
        correct += (outputs.argmax(1) == labels).type(torch.float).sum()
    assert size.item() == 200.0 and abs(correct.item() - ref_correct.item()) <= 2
    e.close()


def test_shim_rejects_misuse():
    from ddp_practice_b200.shim import GradScaler, nn
    model = nn.parallel.DistributedDataParallel(RefConvNet(torch.float16, "cuda").cuda(), device_ids=[0])
    opt = torch.optim.SGD(model.parameters(), 1e-4)
    with pytest.raises(RuntimeError, match="backward"):
        GradScaler().step(opt)
    with pytest.raises(NotImplementedError):
        GradScaler().step(torch.optim.SGD(model.parameters(), 1e-4, maximize=True))


def test_checkpoint_save_load_resume_is_bit_exact(tmp_path):
    """the reference only saves (`ddp_main.py:165-169`); loading the same dict back into a fresh
    model + scaler and continuing must give exactly the run that never stopped.  The file is the reference's format:
    {"model": DDP state_dict with `module.` keys, "scaler": GradScaler.state_dict()}."""
    from ddp_practice_b200.shim import GradScaler, nn

    def build(seed):
        torch.manual_seed(seed)
        m = nn.parallel.DistributedDataParallel(
            nn.SyncBatchNorm.convert_sync_batchnorm(RefConvNet(torch.float16, "cuda").cuda()), device_ids=[0])
        return m, nn.CrossEntropyLoss().cuda(), torch.optim.SGD(m.parameters(), 1e-4), GradScaler(growth_interval=3)

    x, y = O.synthetic_mnist(32 * 8)
    batches = [(x[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32]) for i in range(8)]
    # uninterrupted: 8 steps
    m0, c0, o0, s0 = build(3407)
    _reference_style_loop(m0, batches, c0, o0, s0)
    # interrupted after 4 steps, saved, loaded into objects built from another seed, 4 more steps
    m1, c1, o1, s1 = build(3407)
    _reference_style_loop(m1, batches[:4], c1, o1, s1)
    path = str(tmp_path / "ddp_checkpoint.pt")
    torch.save({"model": m1.state_dict(), "scaler": s1.state_dict()}, path)
    ck = torch.load(path)
    assert set(ck) == {"model", "scaler"} and len(ck["model"]) == 16 and all(k.startswith("module.") for k in ck["model"])
    assert set(ck["scaler"]) == {"scale", "growth_factor", "backoff_factor", "growth_interval", "_growth_tracker"}
    m2, c2, o2, s2 = build(1)
    m2.load_state_dict(ck["model"])
    s2.load_state_dict(ck["scaler"])
    _reference_style_loop(m2, batches[4:], c2, o2, s2)
    torch.cuda.synchronize()
    assert torch.equal(flat_params(m2), flat_params(m0))
    assert torch.equal(bn_buffers(m2), bn_buffers(m0))
    assert s2.state_dict() == s0.state_dict()
    assert m2.state_dict()["module.layer1.1.num_batches_tracked"].item() == 8
    # and the file loads into the plain torch model the reference would use
    ref = RefConvNet(torch.float16, "cuda")
    ref.load_state_dict({k.removeprefix("module."): v.cpu() for k, v in ck["model"].items()})


def test_momentum_sgd_through_the_shim_and_optimizer_checkpoint_resume(tmp_path):
    """SURVEY 8(f) N3 through the reference's call surface: SGD(momentum, weight_decay) handed to scaler.step();
    `optimizer.state_dict()` carries the momentum buffers (views of the library's memory), and a run resumed from
    {"model", "scaler", "optimizer"} continues bit for bit."""
    from ddp_practice_b200.shim import GradScaler, nn
    kw = dict(lr=1e-3, momentum=0.9, weight_decay=1e-3)

    def build(seed):
        torch.manual_seed(seed)
        m = nn.parallel.DistributedDataParallel(
            nn.SyncBatchNorm.convert_sync_batchnorm(RefConvNet(torch.float16, "cuda").cuda()), device_ids=[0])
        return m, nn.CrossEntropyLoss().cuda(), torch.optim.SGD(m.parameters(), **kw), GradScaler()

    x, y = O.synthetic_mnist(32 * 6)
    batches = [(x[i * 32:(i + 1) * 32], y[i * 32:(i + 1) * 32]) for i in range(6)]
    m0, c0, o0, s0 = build(3407)
    _reference_style_loop(m0, batches, c0, o0, s0)
    torch.cuda.synchronize()
    bufs = [o0.state[p]["momentum_buffer"] for p in o0.param_groups[0]["params"]]
    assert torch.equal(torch.cat([b.reshape(-1) for b in bufs]), m0.engine.momentum_buffer)
    assert float(torch.cat([b.reshape(-1) for b in bufs]).abs().sum()) > 0
    # same thing straight through the engine
    from ddp_practice_b200 import Engine
    torch.manual_seed(3407)
    st = O.OracleState.from_state_dict(RefConvNet(torch.float16, "cuda").state_dict())
    e = Engine(dtype="fp16", batch_cap=32, **kw)
    bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
    e.load_state(st.flat_params().numpy(), bn.numpy())
    for xb, yb in batches:
        e.train_step(xb.cuda().contiguous(), yb.cuda().contiguous())
    torch.cuda.synchronize()
    assert torch.equal(e.params.cpu(), flat_params(m0)) and torch.equal(e.momentum_buffer, m0.engine.momentum_buffer)
    e.close()
    # interrupted after 3 steps, resumed in fresh objects
    m1, c1, o1, s1 = build(3407)
    _reference_style_loop(m1, batches[:3], c1, o1, s1)
    path = str(tmp_path / "ck.pt")
    torch.save({"model": m1.state_dict(), "scaler": s1.state_dict(), "optimizer": o1.state_dict()}, path)
    ck = torch.load(path)
    assert len(ck["optimizer"]["state"]) == 10 and ck["optimizer"]["param_groups"][0]["momentum"] == 0.9
    m2, c2, o2, s2 = build(1)
    m2.load_state_dict(ck["model"]); s2.load_state_dict(ck["scaler"]); o2.load_state_dict(ck["optimizer"])
    _reference_style_loop(m2, batches[3:], c2, o2, s2)
    torch.cuda.synchronize()
    assert torch.equal(flat_params(m2), flat_params(m0))
    assert torch.equal(m2.engine.momentum_buffer, m0.engine.momentum_buffer)


def test_shim_lr_scheduler_and_no_sync_keep_one_context():
    """an LR scheduler stepping `param_groups[0]['lr']` must not rebuild the native context nor
    reset the live GradScaler state; `with model.no_sync():` accumulates like torch's DDP.  Loop written the way a user
    of the reference would extend it; compared bit for bit with the C ABI driven directly."""
    from ddp_practice_b200 import Engine
    from ddp_practice_b200.shim import GradScaler, nn
    torch.manual_seed(3407)
    model = RefConvNet(torch.float16, "cuda").cuda()
    init_p, init_bn = flat_params(model).numpy(), bn_buffers(model).numpy()
    model = nn.parallel.DistributedDataParallel(nn.SyncBatchNorm.convert_sync_batchnorm(model), device_ids=[0])
    criterion = nn.CrossEntropyLoss().cuda()
    optimizer = torch.optim.SGD(model.parameters(), 1e-2, momentum=0.9)
    sched = torch.optim.lr_scheduler.StepLR(optimizer, step_size=1, gamma=0.5)
    scaler = GradScaler(growth_interval=3)
    x, y = O.synthetic_mnist(32 * 8)
    x, y = x.cuda(), y.cuda()
    e = Engine(dtype="fp16", batch_cap=32, lr=1e-2, momentum=0.9, growth_interval=3)
    e.load_state(init_p, init_bn)
    model.train()
    engines = set()
    for s in range(4):
        xa, ya, xb, yb = x[64 * s:64 * s + 32], y[64 * s:64 * s + 32], x[64 * s + 32:64 * s + 64], y[64 * s + 32:64 * s + 64]
        optimizer.zero_grad()
        with model.no_sync():  # micro-step: the gradient stays in the accumulation buffer
            scaler.scale(criterion(model(xa), ya)).backward()
        loss = criterion(model(xb), yb)
        scaler.scale(loss).backward()
        scaler.step(optimizer)
        scaler.update()
        sched.step()
        engines.add(id(model.engine))
        e.train_step_accumulate(xa.contiguous(), ya.contiguous())
        e.train_step(xb.contiguous(), yb.contiguous())
        e.set_sgd(optimizer.param_groups[0]["lr"], momentum=0.9)
        torch.cuda.synchronize()
        assert loss.item() == e.loss.item()
        assert scaler.get_scale() == e.scale.item() and scaler.state_dict()["_growth_tracker"] == e.growth_tracker.item()
        assert torch.equal(flat_params(model), e.params.cpu()), f"step {s}"
    assert len(engines) == 1, "the native context was rebuilt"
    assert optimizer.param_groups[0]["lr"] == pytest.approx(1e-2 * 0.5 ** 4)
    e.close()


def test_torch_ops_are_registered_with_the_dispatcher_and_match_the_engine():
    """north_star: "delivered as a PyTorch ... extension behind a thin C-ABI" -- torch.ops.ddp_b200.* (torch.library) over
    the same C ABI: same bits as driving the Engine directly."""
    from ddp_practice_b200 import Engine, torch_ops
    torch.manual_seed(3407)
    m = RefConvNet(torch.float16, "cuda")
    a, b = Engine(dtype="fp16", batch_cap=32), Engine(dtype="fp16", batch_cap=32)
    for e in (a, b):
        e.load_state(flat_params(m).numpy(), bn_buffers(m).numpy())
    h = torch_ops.register(b)
    assert "ddp_b200::train_step" in torch._C._dispatch_get_all_op_names()
    x, y = O.synthetic_mnist(96)
    x, y = x.cuda(), y.cuda()
    for s in range(3):
        a.train_step(x[32 * s:32 * s + 32].contiguous(), y[32 * s:32 * s + 32].contiguous())
        loss = torch.ops.ddp_b200.train_step(x[32 * s:32 * s + 32], y[32 * s:32 * s + 32], h)
    torch.cuda.synchronize()
    assert loss.item() == a.loss.item() and torch.equal(a.params, b.params)
    lg = torch.ops.ddp_b200.eval_batch(x[:32], h)
    assert torch.equal(lg, a.eval_batch(x[:32].contiguous()))
    t = torch.arange(10, dtype=torch.float32, device="cuda")
    assert torch.equal(torch.ops.ddp_b200.allreduce_bucket_(t, h), torch.arange(10, dtype=torch.float32, device="cuda"))
    torch_ops.unregister(h)
    a.close(); b.close()


def test_shim_accepts_several_param_groups():
    """`torch.optim.SGD([{...}, {...}])` on the shim model: same bits as the C ABI with `ddpstep_set_sgd_group`."""
    from ddp_practice_b200 import Engine
    from ddp_practice_b200.shim import GradScaler, nn
    torch.manual_seed(3407)
    model = RefConvNet(torch.float16, "cuda").cuda()
    init_p, init_bn = flat_params(model).numpy(), bn_buffers(model).numpy()
    model = nn.parallel.DistributedDataParallel(nn.SyncBatchNorm.convert_sync_batchnorm(model), device_ids=[0])
    criterion = nn.CrossEntropyLoss().cuda()
    fc = list(model.module.fc.parameters())
    rest = [p for n, p in model.module.named_parameters() if not n.startswith("fc.")]
    optimizer = torch.optim.SGD([{"params": rest, "lr": 1e-2, "momentum": 0.9}, {"params": fc, "lr": 1e-3, "weight_decay": 1e-2}], lr=1e-4)
    scaler = GradScaler()
    e = Engine(dtype="fp16", batch_cap=32)
    e.load_state(init_p, init_bn)
    e.set_sgd_group(0b0011111111, 1e-2, momentum=0.9)
    e.set_sgd_group(0b1100000000, 1e-3, weight_decay=1e-2)
    x, y = O.synthetic_mnist(32 * 4)
    x, y = x.cuda(), y.cuda()
    model.train()
    for s in range(4):
        xb, yb = x[32 * s:32 * s + 32], y[32 * s:32 * s + 32]
        loss = criterion(model(xb), yb)
        optimizer.zero_grad()
        scaler.scale(loss).backward()
        scaler.step(optimizer)
        scaler.update()
        e.train_step(xb.contiguous(), yb.contiguous())
        torch.cuda.synchronize()
        assert torch.equal(flat_params(model), e.params.cpu()), f"step {s}"
    assert "momentum_buffer" in optimizer.state[rest[0]]
    e.close()
