"""Drop-in mirror of the PyTorch call surface used by the reference's DDP scripts.

`ddp_main.py` / `ddp_main_torchrun.py` run on the fused CUDA step with two import edits:

    import torch.nn as nn                    ->  from ddp_practice_b200.shim import nn
    from torch.cuda.amp import GradScaler    ->  from ddp_practice_b200.shim import GradScaler

Everything else in the scripts stays as written (reference call sites cited per class).
`nn` forwards to `torch.nn` except for `SyncBatchNorm`, `parallel.DistributedDataParallel` and
`CrossEntropyLoss`.  The six eager calls of one training iteration (ddp_main.py:88-93) become
lazy handles plus ONE native launch at `scaler.step(optimizer)`.  There is no fallback: a
module that is not the reference's ConvNet layer stack is rejected at wrap time.
"""
from __future__ import annotations

import contextlib
import math
import os
import types
from typing import Optional

import torch
import torch.distributed as dist
import torch.nn as _tnn

from .engine import Engine

_SPECS = [  # (state-dict key, shape) in the library's flat order
    ("layer1.0.weight", (16, 1, 5, 5)), ("layer1.0.bias", (16,)), ("layer1.1.weight", (16,)), ("layer1.1.bias", (16,)),
    ("layer2.0.weight", (32, 16, 5, 5)), ("layer2.0.bias", (32,)), ("layer2.1.weight", (32,)), ("layer2.1.bias", (32,)),
    ("fc.weight", (10, 1568)), ("fc.bias", (10,)),
]
_BUFS = [("layer1.1.running_mean", 0, 16), ("layer1.1.running_var", 16, 16), ("layer2.1.running_mean", 32, 32),
         ("layer2.1.running_var", 64, 32)]
_DT_NAME = {torch.float32: "fp32", torch.float16: "fp16", torch.bfloat16: "bf16"}


def _check_convnet(m: _tnn.Module):
    """Accept exactly the layer stack of ddp_main.py:13-28; anything else is an error."""
    def block(seq, cin, cout):
        ok = isinstance(seq, _tnn.Sequential) and len(seq) == 4
        ok = ok and isinstance(seq[0], _tnn.Conv2d) and (seq[0].in_channels, seq[0].out_channels) == (cin, cout)
        ok = ok and seq[0].kernel_size == (5, 5) and seq[0].stride == (1, 1) and seq[0].padding == (2, 2)
        ok = ok and seq[0].bias is not None and seq[0].groups == 1 and seq[0].dilation == (1, 1)
        ok = ok and isinstance(seq[1], _tnn.modules.batchnorm._BatchNorm) and seq[1].num_features == cout
        ok = ok and seq[1].affine and seq[1].track_running_stats and seq[1].momentum == 0.1 and seq[1].eps == 1e-5
        ok = ok and isinstance(seq[2], _tnn.ReLU) and isinstance(seq[3], _tnn.MaxPool2d)
        ok = ok and seq[3].kernel_size in (2, (2, 2)) and seq[3].stride in (2, (2, 2)) and seq[3].padding in (0, (0, 0))
        return ok
    ok = hasattr(m, "layer1") and hasattr(m, "layer2") and hasattr(m, "fc")
    ok = ok and block(m.layer1, 1, 16) and block(m.layer2, 16, 32)
    ok = ok and isinstance(m.fc, _tnn.Linear) and (m.fc.in_features, m.fc.out_features) == (1568, 10)
    ok = ok and m.fc.bias is not None and len(list(m.parameters())) == 10
    if not ok:
        raise NotImplementedError(
            "ddp_practice_b200 accelerates exactly the DDP-practice ConvNet ([Conv5x5 p2, BN, ReLU, MaxPool2] x2 + "
            "Linear(1568,10)); this module is something else and there is no fallback path")


class _Pending:
    """What `model(images)` returns in training mode: the batch, not yet computed."""

    def __init__(self, owner, images):
        self.owner, self.images = owner, images

    def _materialised(self):
        raise RuntimeError("training-mode outputs are produced inside the fused step; read them from "
                           "`model.last_logits` after scaler.step(optimizer)")

    argmax = size = float = __getitem__ = lambda self, *a, **k: self._materialised()


class _Loss:
    """What `criterion(outputs, labels)` returns: usable by scaler.scale() and, after the step, .item()."""

    def __init__(self, owner, images, labels):
        self.owner, self.images, self.labels = owner, images, labels
        self.step_id = None

    def backward(self):  # origin_main.py:64 style (no scaler)
        self.owner._arm(self, scaled=False)

    def item(self):
        if self.step_id is None or self.owner._steps != self.step_id:
            raise RuntimeError("the loss of a fused step is available after scaler.step(optimizer) / optimizer.step()")
        return float(self.owner.engine.loss.item())

    def __float__(self):
        return self.item()


class _ScaledLoss:
    def __init__(self, loss: _Loss, scaler=None):
        self.loss, self.scaler = loss, scaler

    def backward(self):  # ddp_main.py:91
        self.loss.owner._arm(self.loss, scaled=True, scaler=self.scaler)


class DistributedDataParallel(_tnn.Module):
    """`nn.parallel.DistributedDataParallel(model, device_ids=[local_rank])`, ddp_main.py:121-123.

    Construction does what DDP's does that is observable: rank 0's parameters and buffers are
    broadcast (distributed.py:881-890).  The native context is created at the first step, when
    the scaler settings and learning rate are known."""

    def __init__(self, module: _tnn.Module, device_ids=None, output_device=None, **unsupported):
        super().__init__()
        for k, v in unsupported.items():
            if v not in (None, False, 25):  # defaults of find_unused_parameters / bucket_cap_mb etc.
                raise NotImplementedError(f"DistributedDataParallel({k}={v!r}) is not supported by the fused step")
        _check_convnet(module)
        p0 = next(module.parameters())
        if not p0.is_cuda:
            raise RuntimeError("the module must be on a CUDA device (ConvNet().cuda()); there is no CPU path")
        self.module = module
        self.device_index = p0.device.index if device_ids is None else int(device_ids[0])
        self.world = dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank() if self.world > 1 else 0
        self.sync_bn = bool(getattr(module, "_b200_sync_bn", False)) or any(
            isinstance(mm, _tnn.SyncBatchNorm) for mm in module.modules())
        if self.world > 1:
            with torch.no_grad():
                for t in list(module.parameters()) + list(module.buffers()):
                    dist.broadcast(t.data, 0)
        self.amp_dtype = self._probe_dtype()
        self.batch_cap = 0
        self._scaler_cfg = None
        self.engine: Optional[Engine] = None
        self._engine_key = None
        self._armed = None
        self._steps = 0
        self._mom_published = False  # optimizer.state[p]['momentum_buffer'] views handed out
        self._no_sync = False
        self._groups_key = None      # per-tensor settings last sent for an optimizer with several param groups
        self._optimizer = None       # the optimizer seen by the last step (its state holds views of our memory)
        self.last_logits = None

    # -- construction helpers -------------------------------------------------------------
    def _probe_dtype(self) -> str:
        """Which dtype does the user's forward compute in?  One tiny eval-mode forward of the
        user's own module at construction time (never on the hot path): fp16 for
        `torch.cuda.amp.autocast()` (ddp_main.py:31), bf16 for autocast(dtype=bfloat16), fp32
        for origin_main.py's forward."""
        was = self.module.training
        self.module.eval()
        try:
            with torch.no_grad():
                out = self.module(torch.zeros(1, 1, 28, 28, device=next(self.module.parameters()).device))
        finally:
            self.module.train(was)
        return _DT_NAME[out.dtype]

    def _named(self):
        sd_p = dict(self.module.named_parameters())
        sd_b = dict(self.module.named_buffers())
        return sd_p, sd_b

    def _ensure_engine(self, use_scaler: bool, lr: float = 1e-4, scaler: "GradScaler" = None, sgd: dict = None,
                       batch: int = 0):
        """The native context is built at the first step (or first eval).  What may change afterwards -- lr and the
        other SGD knobs of `optimizer.param_groups[0]`, e.g. under an LR scheduler -- is run-time state of the
        library (`ddpstep_set_sgd`): no new context, no IPC re-exchange, no graph re-capture, and the live
        GradScaler state is never touched.  Only a batch larger than any seen so far, or a GradScaler whose
        constructor arguments differ from what an eval-first context assumed, rebuild it -- carrying every piece of
        live state (parameters, BN buffers, scale, growth tracker, momentum buffers) over."""
        sgd = sgd or {}
        hyper = (float(lr), tuple(sorted(sgd.items())))
        scfg = None if scaler is None else (scaler._growth_factor, scaler._backoff_factor, scaler._growth_interval)
        carry = None
        if self.engine is not None:
            if self._steps > 0 and use_scaler != self._engine_key[0]:
                raise RuntimeError("one model cannot mix scaler.step(optimizer) and optimizer.step(): the native "
                                   "context is built for one of the two (ddp_main.py:91-93 vs origin_main.py:64-66)")
            rebuild = batch > self.engine.batch_cap or use_scaler != self._engine_key[0]
            rebuild = rebuild or (scfg is not None and self._scaler_cfg is not None and scfg != self._scaler_cfg)
            if not rebuild:
                if scfg is not None and self._scaler_cfg is None:  # eval-first context, scaler defaults matched
                    self._scaler_cfg = scfg
                    if scaler._loaded or scaler._init_scale != 65536.0:
                        self.engine.set_scaler(scaler._init_scale_now(), scaler._tracker_now())
                if hyper != self._engine_key[1]:
                    self.engine.set_sgd(float(lr), float(sgd.get("momentum", 0.0)), float(sgd.get("weight_decay", 0.0)),
                                        float(sgd.get("dampening", 0.0)), bool(sgd.get("nesterov", False)))
                    self._engine_key = (use_scaler, hyper)
                return self.engine
            old = self.engine
            old.check_device()
            carry = {"scale": float(old.scale.item()), "tracker": int(old.growth_tracker.item()),
                     "mom": old.momentum_buffer.clone() if int(old.momentum_ready.item()) else None,
                     "stepped": self._steps > 0}
            opt = self._optimizer
            if opt is not None:  # its momentum_buffer entries are views of the context about to be freed
                lo = old.momentum_buffer.data_ptr()
                for st in opt.state.values():
                    buf = st.get("momentum_buffer")
                    if buf is not None and lo <= buf.data_ptr() < lo + 4 * old.momentum_buffer.numel():
                        del st["momentum_buffer"]
        key = (use_scaler, hyper)
        sd_p, sd_b = self._named()
        flat = torch.cat([sd_p[n].detach().reshape(-1).float() for n, _ in _SPECS]).cpu()
        bn = torch.cat([sd_b[n].detach().float() for n, _, _ in _BUFS]).cpu()
        nbt = [int(sd_b["layer1.1.num_batches_tracked"]), int(sd_b["layer2.1.num_batches_tracked"])]
        if carry is not None:
            self.engine.close()
            self.engine = None
        kw = {}
        if scaler is not None:
            kw = dict(init_scale=scaler._init_scale_now(), growth_factor=scaler._growth_factor,
                      backoff_factor=scaler._backoff_factor, growth_interval=scaler._growth_interval)
        tracker = scaler._tracker_now() if scaler is not None else 0
        # sized for the batches actually seen (ddp_main.py:51 `-b`, default 32), not for the largest one imaginable;
        # DDP_B200_BATCH_CAP pre-sizes it
        cap = max(int(os.environ.get("DDP_B200_BATCH_CAP", "0")), int(batch), self.batch_cap, 32)
        self.batch_cap = cap
        eng = Engine(dtype=self.amp_dtype, batch_cap=cap, rank=self.rank, world=self.world,
                     device=self.device_index, use_scaler=use_scaler, sync_bn=self.sync_bn, lr=lr, **kw, **sgd)
        eng.connect()
        eng.load_state(flat.numpy(), bn.numpy(), nbt)
        if carry is not None and carry["stepped"] and use_scaler:
            eng.set_scaler(carry["scale"], carry["tracker"])  # the LIVE scaler state, not the constructor's
        elif scaler is not None and tracker:
            eng.set_scaler(kw["init_scale"], tracker)
        if carry is not None and carry["mom"] is not None:
            eng.momentum_buffer.copy_(carry["mom"])
            eng.momentum_ready.fill_(1)
        self._scaler_cfg = scfg
        # re-point the module's tensors at the library's memory: state_dict(), parameters(),
        # torch.save and p.grad keep working and always show the live values
        off = 0
        for name, shape in _SPECS:
            n = math.prod(shape)
            p = sd_p[name]
            p.data = eng.params[off:off + n].view(shape)
            p._b200_owner = self
            p._b200_slice = (off, n, shape)
            off += n
        for name, o, n in _BUFS:
            sd_b[name].data = eng.bn_buffers[o:o + n]
        sd_b["layer1.1.num_batches_tracked"].data = eng.num_batches_tracked[0]
        sd_b["layer2.1.num_batches_tracked"].data = eng.num_batches_tracked[1]
        self.engine, self._engine_key = eng, key
        self._groups_key = None  # a new context starts with the first group's settings everywhere: re-send the others
        if carry is not None:
            self._mom_published = False  # optimizer.state views point into the old context: re-publish
        return eng

    # -- failure reporting ----------------------------------------------------------------
    def state_dict(self, *a, **k):
        """`model.state_dict()` (ddp_main.py:166): the tensors alias the library's memory, so first make sure no
        step was voided by a peer timeout (raises NativeError / DDPSTEP_E_PEER otherwise)."""
        if self.engine is not None:
            self.engine.check_device()
        return super().state_dict(*a, **k)

    def train(self, mode: bool = True):
        if self.engine is not None and not mode:  # entering test() (ddp_main.py:98): once per epoch
            self.engine.check_device()
        return super().train(mode)

    def _attach_grads(self):
        sd_p, _ = self._named()
        for name, _shape in _SPECS:
            p = sd_p[name]
            off, n, shape = p._b200_slice
            p.grad = self.engine.grads[off:off + n].view(shape)

    # -- the call surface -----------------------------------------------------------------
    def forward(self, images):
        if images.dim() != 4 or tuple(images.shape[1:]) != (1, 28, 28):
            raise ValueError(f"expected [B,1,28,28] images, got {tuple(images.shape)}")
        images = images.contiguous().float()
        if self.training:
            return _Pending(self, images)  # ddp_main.py:88
        eng = self.engine
        if eng is None or images.shape[0] > eng.batch_cap:
            key = self._engine_key or (self.amp_dtype != "fp32", (1e-4, ()))
            eng = self._ensure_engine(key[0], key[1][0], None, dict(key[1][1]), batch=images.shape[0])
        return eng.eval_batch(images)  # ddp_main.py:104: real logits tensor

    def _arm(self, loss: _Loss, scaled: bool, scaler=None):
        """`.backward()`.  Normally this only arms the step that scaler.step(optimizer) / optimizer.step() launches.
        Inside `with model.no_sync():` (gradient accumulation, distributed.py:1441-1470) there is no optimizer step to
        wait for: the micro-step runs now and leaves its gradient in the library's accumulation buffer."""
        if scaler is not None:
            scaler._owner = self
        if self._no_sync:
            key = self._engine_key or (scaled, (1e-4, ()))
            eng = self._ensure_engine(scaled, key[1][0], scaler, dict(key[1][1]), batch=loss.images.shape[0])
            labels = loss.labels.contiguous()
            eng.train_step_accumulate(loss.images, labels if labels.dtype == torch.int64 else labels.long())
            self._steps += 1
            loss.step_id = self._steps
            self.last_logits = eng.logits[: loss.images.shape[0]]
            return
        self._armed = (loss, scaled)

    @contextlib.contextmanager
    def no_sync(self):
        """`with model.no_sync():` -- backward passes inside accumulate locally; the first one outside all-reduces
        the accumulated gradient together with its own (same contract as torch's DDP.no_sync)."""
        old, self._no_sync = self._no_sync, True
        try:
            yield
        finally:
            self._no_sync = old

    def _apply_param_groups(self, optimizer):
        """`optimizer.param_groups` with different settings (torch/optim/sgd.py runs its update once per group): every
        group becomes one `ddpstep_set_sgd_group` call naming its parameter tensors; re-sent only when something changed."""
        index = {id(p): i for i, (name, _) in enumerate(_SPECS) for p in [dict(self.module.named_parameters())[name]]}
        groups, seen = [], 0
        for pg in optimizer.param_groups:
            mask = 0
            for p in pg["params"]:
                if id(p) not in index:
                    raise RuntimeError("the optimizer holds a parameter that is not one of this model's")
                mask |= 1 << index[id(p)]
            seen |= mask
            k = _sgd_knobs(pg)
            groups.append((mask, float(pg["lr"]), k["momentum"], k["weight_decay"], k["dampening"], k["nesterov"]))
        if seen != 0x3FF:
            raise NotImplementedError("every parameter of the model must be in some param group: torch would leave the "
                                      "others un-updated, the fused tail updates all ten tensors")
        key = tuple(groups)
        if key != self._groups_key:
            for mask, lr, mom, wd, damp, nest in groups:
                self.engine.set_sgd_group(mask, lr, mom, wd, damp, nest)
            self._groups_key = key

    def _sync_momentum(self, optimizer, before: bool):
        """`optimizer.state[p]['momentum_buffer']` are views of the library's flat buffer.  Before a step: adopt
        buffers that came from `optimizer.load_state_dict`; after the first executed step: publish the views."""
        eng = self.engine
        sd_p, _ = self._named()
        for name, _shape in _SPECS:
            p = sd_p[name]
            off, n, shape = p._b200_slice
            view = eng.momentum_buffer[off:off + n].view(shape)
            buf = optimizer.state.get(p, {}).get("momentum_buffer")
            if before:
                if buf is not None and buf.data_ptr() != view.data_ptr():
                    view.copy_(buf.to(view.device, torch.float32))
                    eng.momentum_ready.fill_(1)
                    optimizer.state[p]["momentum_buffer"] = view
            elif buf is None and int(eng.momentum_ready.item()):
                optimizer.state[p]["momentum_buffer"] = view

    def _run_step(self, use_scaler: bool, lr: float, scaler=None, sgd: dict = None, optimizer=None):
        if self._armed is None:
            raise RuntimeError("step() without a preceding loss.backward() / scaler.scale(loss).backward()")
        loss, scaled = self._armed
        if scaled != use_scaler:
            raise RuntimeError("scaler.scale(loss).backward() must be followed by scaler.step(optimizer), and "
                               "loss.backward() by optimizer.step()")
        self._optimizer = optimizer
        eng = self._ensure_engine(use_scaler, lr, scaler, sgd, batch=loss.images.shape[0])
        if optimizer is not None and (len(optimizer.param_groups) > 1 or self._groups_key is not None):
            self._apply_param_groups(optimizer)
        momentum = optimizer is not None and any(float(pg.get("momentum", 0.0)) != 0.0 for pg in optimizer.param_groups)
        if momentum and optimizer.state:
            # cheap sentinel: only a load_state_dict() since the last step makes the first buffer something else
            # than our view of the library's memory
            p0 = optimizer.param_groups[0]["params"][0]
            b0 = optimizer.state.get(p0, {}).get("momentum_buffer")
            if b0 is not None and b0.data_ptr() != eng.momentum_buffer.data_ptr():
                self._sync_momentum(optimizer, before=True)
                self._mom_published = True
        labels = loss.labels.contiguous()
        if labels.dtype != torch.int64:
            labels = labels.long()
        eng.train_step(loss.images, labels)
        if momentum and not self._mom_published and int(eng.momentum_ready.item()):
            self._sync_momentum(optimizer, before=False)
            self._mom_published = True
        self._armed = None
        self._steps += 1
        loss.step_id = self._steps
        self.last_logits = eng.logits[: loss.images.shape[0]]
        if os.environ.get("DDP_B200_EXPOSE_GRADS", "1") == "1":
            self._attach_grads()


class CrossEntropyLoss(_tnn.CrossEntropyLoss):
    """`nn.CrossEntropyLoss().cuda()`, ddp_main.py:124,89."""

    def forward(self, outputs, labels):
        if isinstance(outputs, _Pending):
            if self.weight is not None or self.reduction != "mean" or self.label_smoothing != 0.0 or self.ignore_index != -100:
                raise NotImplementedError("the fused step implements CrossEntropyLoss() with default arguments only")
            return _Loss(outputs.owner, outputs.images, labels)
        return super().forward(outputs, labels)


class SyncBatchNorm(_tnn.SyncBatchNorm):
    @classmethod
    def convert_sync_batchnorm(cls, module, process_group=None):
        """`nn.SyncBatchNorm.convert_sync_batchnorm(model)`, ddp_main.py:120: selects BN statistics
        over the global batch; the layers themselves are taken over by the fused step."""
        module._b200_sync_bn = True
        return module


class GradScaler:
    """`GradScaler()`, ddp_main.py:126; defaults of torch/amp/grad_scaler.py:126-129."""

    def __init__(self, device="cuda", init_scale=2.0 ** 16, growth_factor=2.0, backoff_factor=0.5, growth_interval=2000,
                 enabled=True):
        if not enabled:
            raise NotImplementedError("GradScaler(enabled=False): use loss.backward(); optimizer.step() instead")
        self._init_scale, self._growth_factor = float(init_scale), float(growth_factor)
        self._backoff_factor, self._growth_interval = float(backoff_factor), int(growth_interval)
        self._owner: Optional[DistributedDataParallel] = None
        self._loaded = None  # state_dict loaded before the engine exists

    # engine-backed state -----------------------------------------------------------------
    def _eng(self):
        return self._owner.engine if self._owner is not None else None

    def _init_scale_now(self):
        return float(self._loaded["scale"]) if self._loaded else self._init_scale

    def _tracker_now(self):
        return int(self._loaded["_growth_tracker"]) if self._loaded else 0

    @property
    def _scale(self):
        e = self._eng()
        return e.scale if e is not None else None

    @property
    def _growth_tracker(self):
        e = self._eng()
        return e.growth_tracker if e is not None else None

    def get_scale(self):
        e = self._eng()
        return float(e.scale.item()) if e is not None else self._init_scale_now()

    # the three calls of ddp_main.py:91-93 ---------------------------------------------------
    def scale(self, loss):
        if not isinstance(loss, _Loss):
            raise TypeError("scaler.scale() expects the loss returned by the shim's CrossEntropyLoss on model outputs")
        return _ScaledLoss(loss, self)

    def step(self, optimizer):
        params = optimizer.param_groups[0]["params"]
        g = optimizer.param_groups[0]
        if any(pg.get("maximize", False) for pg in optimizer.param_groups):
            raise NotImplementedError("the fused step implements torch.optim.SGD with maximize=False")
        owner = getattr(params[0], "_b200_owner", None)
        if owner is None:  # engine not built yet: find the wrapper through the armed loss
            owner = _find_owner(params)
        self._owner = owner
        owner._run_step(True, float(g["lr"]), self, _sgd_knobs(g), optimizer)
        self._loaded = None

    def update(self, new_scale=None):
        if new_scale is not None:
            raise NotImplementedError("scaler.update(new_scale) is not supported")
        # `_amp_update_scale_` already ran on the device inside the step launch

    def state_dict(self):
        e = self._eng()
        scale = float(e.scale.item()) if e is not None else self._init_scale_now()
        tracker = int(e.growth_tracker.item()) if e is not None else self._tracker_now()
        return {"scale": scale, "growth_factor": self._growth_factor, "backoff_factor": self._backoff_factor,
                "growth_interval": self._growth_interval, "_growth_tracker": tracker}

    def load_state_dict(self, sd):
        self._growth_factor, self._backoff_factor = float(sd["growth_factor"]), float(sd["backoff_factor"])
        self._growth_interval = int(sd["growth_interval"])
        e = self._eng()
        if e is not None:
            e.set_scaler(float(sd["scale"]), int(sd["_growth_tracker"]))
        else:
            self._loaded = dict(sd)


def _sgd_knobs(g) -> dict:
    """momentum / weight_decay / dampening / nesterov of a torch.optim.SGD param group (sgd.py:425-465)."""
    return dict(momentum=float(g.get("momentum", 0.0)), weight_decay=float(g.get("weight_decay", 0.0)),
                dampening=float(g.get("dampening", 0.0)), nesterov=bool(g.get("nesterov", False)))


_OWNERS = []  # wrappers alive in this process, to route an optimizer to its wrapper before the engine exists


def _find_owner(params):
    ids = {id(p) for p in params}
    for o in _OWNERS:
        if any(id(p) in ids for p in o.module.parameters()):
            return o
    raise RuntimeError("this optimizer does not hold the parameters of a shim DistributedDataParallel model")


_orig_init = DistributedDataParallel.__init__


def _init_and_register(self, *a, **k):
    _orig_init(self, *a, **k)
    _OWNERS.append(self)


DistributedDataParallel.__init__ = _init_and_register


class SGD(torch.optim.SGD):
    """`torch.optim.SGD(model.parameters(), 1e-4)` for the no-scaler form (origin_main.py:64-66):
    `loss.backward(); optimizer.step()` runs the fused step without GradScaler."""

    def zero_grad(self, set_to_none=True):
        pass

    def step(self, closure=None):
        g = self.param_groups[0]
        owner = getattr(g["params"][0], "_b200_owner", None) or _find_owner(g["params"])
        owner._run_step(False, float(g["lr"]), None, _sgd_knobs(g), self)


class _NN(types.ModuleType):
    """`nn`: torch.nn with the three names the fused step takes over."""

    def __getattr__(self, name):
        return getattr(_tnn, name)


nn = _NN("ddp_practice_b200.shim.nn")
nn.SyncBatchNorm = SyncBatchNorm
nn.CrossEntropyLoss = CrossEntropyLoss
nn.parallel = types.SimpleNamespace(DistributedDataParallel=DistributedDataParallel)
