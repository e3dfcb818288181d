"""CPU oracle for the DDP+AMP ConvNet training step.  TEST INFRASTRUCTURE ONLY.

This file is the checker, never the product: only `tests/`,
`__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py` may import it.  The shipped path (`ddp_practice_b200`) never does and
fails loudly when its CUDA library is missing.

What it restates (reference = gbbin/DDP-practice, arithmetic lives in torch):

* the ConvNet layer stack                         ddp_main.py:13-36, origin_main.py:9-31
* one `train()` iteration                         ddp_main.py:85-93, origin_main.py:60-66
* one `test()` iteration + the two reductions     ddp_main.py:96-112
* autocast rounding points                        torch/include/ATen/autocast_mode.h:820-853
* SyncBatchNorm forward/backward algebra          torch/nn/modules/_functions.py:10-209
* DDP Reducer: pre-divide by W then SUM           torch/.../c10d/reducer.hpp:276, default_comm_hooks.hpp:36-51
* GradScaler scale/unscale_/step/update           torch/amp/grad_scaler.py:193-217,302-373,484-559
* SGD (momentum 0, wd 0) foreach update           torch/optim/sgd.py:473
* gradient accumulation under DDP.no_sync()       torch/nn/parallel/distributed.py:1441-1470

The contractions (conv2d / matmul) are delegated to ATen's fp32 CPU kernels on
operands that were first rounded to the autocast dtype; every other formula
(BN statistics, merge, ReLU/pool routing, CE, scaler state machine) is spelled
out.  W ranks are simulated in one process: each rank runs its own shard, the
collectives become plain Python sums in rank order.

Parity pinning: the reference ships no tests or golden vectors, so this oracle
is pinned against the reference code itself, imported from /root/reference in
the build container by `tests/golden/make_golden.py`; the resulting vectors are
committed under `tests/golden/` and checked by `tests/test_oracle_golden.py`.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

# parameter layout shared with the CUDA library (definition order, flat fp32)
PARAM_SPECS = [
    ("layer1.0.weight", (16, 1, 5, 5)),
    ("layer1.0.bias", (16,)),
    ("layer1.1.weight", (16,)),
    ("layer1.1.bias", (16,)),
    ("layer2.0.weight", (32, 16, 5, 5)),
    ("layer2.0.bias", (32,)),
    ("layer2.1.weight", (32,)),
    ("layer2.1.bias", (32,)),
    ("fc.weight", (10, 1568)),
    ("fc.bias", (10,)),
]
N_PARAMS = sum(math.prod(s) for _, s in PARAM_SPECS)  # 29,034
BN_EPS = 1e-5
BN_MOMENTUM = 0.1

_DT = {"fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16}


def rnd(t: torch.Tensor, dt: torch.dtype) -> torch.Tensor:
    """Round an fp32 tensor to the autocast storage dtype and come back to fp32."""
    if dt == torch.float32:
        return t
    return t.to(dt).to(torch.float32)


@dataclass
class OracleState:
    """fp32 master parameters, BN buffers and GradScaler state of one replica."""

    params: Dict[str, torch.Tensor]
    running_mean: List[torch.Tensor]
    running_var: List[torch.Tensor]
    num_batches_tracked: List[int] = field(default_factory=lambda: [0, 0])
    scale: float = 65536.0  # grad_scaler.py:126
    growth_tracker: int = 0
    growth_factor: float = 2.0
    backoff_factor: float = 0.5
    growth_interval: int = 2000
    # torch.optim.SGD state (sgd.py:425-465): flat momentum buffer, created by the first executed step
    momentum_buffer: Optional[torch.Tensor] = None
    # DDP.no_sync() / gradient accumulation (torch/nn/parallel/distributed.py:1441-1470): what the `.grad` tensors hold
    # between two optimizer steps when zero_grad() is not called -- one flat fp32 gradient per rank, None when empty
    grad_accum: Optional[List[torch.Tensor]] = None

    @staticmethod
    def from_state_dict(sd: Dict[str, torch.Tensor]) -> "OracleState":
        sd = {k.removeprefix("module."): v for k, v in sd.items()}
        return OracleState(
            params={n: sd[n].detach().clone().float().reshape(s) for n, s in PARAM_SPECS},
            running_mean=[sd["layer1.1.running_mean"].clone().float(), sd["layer2.1.running_mean"].clone().float()],
            running_var=[sd["layer1.1.running_var"].clone().float(), sd["layer2.1.running_var"].clone().float()],
            num_batches_tracked=[int(sd["layer1.1.num_batches_tracked"]), int(sd["layer2.1.num_batches_tracked"])],
        )

    def flat_params(self) -> torch.Tensor:
        return torch.cat([self.params[n].reshape(-1) for n, _ in PARAM_SPECS])

    def clone(self) -> "OracleState":
        return OracleState(
            params={k: v.clone() for k, v in self.params.items()},
            running_mean=[t.clone() for t in self.running_mean],
            running_var=[t.clone() for t in self.running_var],
            num_batches_tracked=list(self.num_batches_tracked),
            scale=self.scale, growth_tracker=self.growth_tracker,
            growth_factor=self.growth_factor, backoff_factor=self.backoff_factor,
            growth_interval=self.growth_interval,
            momentum_buffer=None if self.momentum_buffer is None else self.momentum_buffer.clone(),
        )


def flatten_grads(grads: Dict[str, torch.Tensor]) -> torch.Tensor:
    return torch.cat([grads[n].reshape(-1) for n, _ in PARAM_SPECS])


# --------------------------------------------------------------------------- forward pieces


# How the conv bias meets the rounding.  "double": ATen's cuDNN path runs
# `cudnn_convolution` (output rounded to `dt`) and then `output.add_(bias)` (rounded
# again) -- what the reference does on a GPU.  "single": bias folded into the fp32
# accumulator and rounded once -- what ATen's CPU (oneDNN) kernel does, and therefore
# what the CPU-generated golden vectors in tests/golden/ contain.
CONV_BIAS_ROUNDING = "double"


def _conv(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, dt: torch.dtype) -> torch.Tensor:
    """Conv2d 5x5 pad 2 under autocast: operands in `dt`, fp32 accumulate, `dt` output."""
    acc = F.conv2d(x, rnd(w, dt), None, 1, 2)
    if CONV_BIAS_ROUNDING == "double":
        acc = rnd(acc, dt)
    return rnd(acc + rnd(b, dt).view(1, -1, 1, 1), dt)


def _local_stats(y: torch.Tensor):
    """`batch_norm_stats`: per-channel mean and 1/sqrt(biased var + eps), fp32."""
    n = y.numel() // y.shape[1]
    mean = y.mean(dim=(0, 2, 3))
    var = ((y - mean.view(1, -1, 1, 1)) ** 2).mean(dim=(0, 2, 3))
    return mean, 1.0 / torch.sqrt(var + BN_EPS), n


def merge_stats(means: Sequence[torch.Tensor], invstds: Sequence[torch.Tensor], counts: Sequence[int]):
    """`batch_norm_gather_stats_with_counts` (_functions.py:106): Chan merge in rank
    order, fp32.  Returns global mean, invstd, biased var·n (M2) and n."""
    mean = torch.zeros_like(means[0])
    m2 = torch.zeros_like(means[0])
    n = torch.zeros((), dtype=torch.float32)
    for mu_r, is_r, n_r in zip(means, invstds, counts):
        n_r = torch.tensor(float(n_r), dtype=torch.float32)
        v = (1.0 / (is_r * is_r) - BN_EPS) * n_r
        f = 1.0 / (n + n_r)
        m2 = m2 + v + (mean - mu_r) ** 2 * n * n_r * f
        mean = n * f * mean + n_r * f * mu_r
        n = n + n_r
    invstd = 1.0 / torch.sqrt(m2 / n + BN_EPS)
    return mean, invstd, m2, float(n)


def _bn_apply(y, mean, invstd, gamma, beta, dt):
    c = lambda t: t.view(1, -1, 1, 1)
    return rnd((y - c(mean)) * c(invstd) * c(gamma) + c(beta), dt)


def _relu_pool(a: torch.Tensor):
    """ReLU then MaxPool2d(2,2); ties go to the first element in scan order
    (top-left, top-right, bottom-left, bottom-right).  Returns pooled values and
    the 0..3 window position that won."""
    r = torch.clamp_min(a, 0.0)
    B, C, H, W = r.shape
    win = r.view(B, C, H // 2, 2, W // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, H // 2, W // 2, 4)
    arg = win.argmax(dim=-1)  # first maximum
    return win.gather(-1, arg.unsqueeze(-1)).squeeze(-1), arg


def _unpool_relu_bwd(dp: torch.Tensor, arg: torch.Tensor, a: torch.Tensor) -> torch.Tensor:
    """max_pool2d backward (route to the winner) followed by threshold_backward
    (gradient passes where the ReLU output is > 0)."""
    B, C, h, w = dp.shape
    win = torch.zeros(B, C, h, w, 4, dtype=dp.dtype)
    win.scatter_(-1, arg.unsqueeze(-1), dp.unsqueeze(-1))
    d = win.view(B, C, h, w, 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(B, C, 2 * h, 2 * w)
    # `where`, not a multiply: a non-finite upstream gradient must stay confined
    # to the winner (inf * 0 would poison the whole window with NaN)
    return torch.where(a > 0, d, torch.zeros_like(d))


# --------------------------------------------------------------------------- one rank, forward + backward


class _RankPass:
    """One rank's forward/backward over its shard, with the SyncBN exchange points
    exposed so that `train_step` can play the collectives."""

    def __init__(self, st: OracleState, x: torch.Tensor, y: torch.Tensor, dt: torch.dtype):
        self.st, self.dt = st, dt
        self.x = rnd(x.float(), dt)
        self.y = y.long()
        self.B = x.shape[0]

    # forward, stage by stage ------------------------------------------------
    def conv1(self):
        p = self.st.params
        self.y1 = _conv(self.x, p["layer1.0.weight"], p["layer1.0.bias"], self.dt)
        return _local_stats(self.y1)

    def bn1_conv2(self, mean, invstd):
        p = self.st.params
        self.mu1, self.is1 = mean, invstd
        self.a1 = _bn_apply(self.y1, mean, invstd, p["layer1.1.weight"], p["layer1.1.bias"], self.dt)
        self.p1, self.arg1 = _relu_pool(self.a1)
        self.y2 = _conv(self.p1, p["layer2.0.weight"], p["layer2.0.bias"], self.dt)
        return _local_stats(self.y2)

    def bn2_head(self, mean, invstd):
        p, dt = self.st.params, self.dt
        self.mu2, self.is2 = mean, invstd
        self.a2 = _bn_apply(self.y2, mean, invstd, p["layer2.1.weight"], p["layer2.1.bias"], dt)
        self.p2, self.arg2 = _relu_pool(self.a2)
        self.flat = self.p2.reshape(self.B, -1)
        self.logits = rnd(self.flat @ rnd(p["fc.weight"], dt).t() + rnd(p["fc.bias"], dt), dt)
        # CrossEntropyLoss outside autocast on `dt` logits: fp32 inside, `dt` results
        self.lsm = rnd(torch.log_softmax(self.logits, dim=1), dt)
        nll = -self.lsm.gather(1, self.y.view(-1, 1)).squeeze(1)
        self.loss = rnd(nll.mean(), dt)
        return self.loss

    # backward ----------------------------------------------------------------
    def head_bwd(self, scale: float):
        """d(loss*scale): the incoming gradient is `scale` cast to the loss dtype
        (65536 -> inf in fp16, grad_scaler.py:217 + autograd cast)."""
        p, dt, B = self.st.params, self.dt, self.B
        gl = rnd(torch.tensor(scale, dtype=torch.float32), dt)
        g = rnd(-(gl / B), dt)
        onehot = F.one_hot(self.y, 10).to(torch.float32)
        gfull = torch.where(onehot > 0, g, torch.zeros(()))
        self.dlogits = rnd(gfull - torch.exp(self.lsm) * gfull.sum(dim=1, keepdim=True), dt)
        self.g = {}
        self.g["fc.weight"] = rnd(self.dlogits.t() @ self.flat, dt)
        self.g["fc.bias"] = rnd(self.dlogits.sum(dim=0), dt)
        dflat = rnd(self.dlogits @ rnd(p["fc.weight"], dt), dt)
        self.da2 = _unpool_relu_bwd(dflat.view(self.B, 32, 7, 7), self.arg2, self.a2)
        return self._bn_bwd_sums(self.da2, self.y2, self.mu2)

    @staticmethod
    def _bn_bwd_sums(da, y, mu):
        """`batch_norm_backward_reduce`: sum(dy) and sum(dy*(x-mean)), fp32."""
        return da.sum(dim=(0, 2, 3)), (da * (y - mu.view(1, -1, 1, 1))).sum(dim=(0, 2, 3))

    def _bn_bwd_elemt(self, da, y, mu, invstd, gamma, sum_dy, sum_dy_xmu, n_total):
        c = lambda t: t.view(1, -1, 1, 1)
        dx = (da - c(sum_dy) / n_total - (y - c(mu)) * c(invstd * invstd * sum_dy_xmu / n_total)) * c(invstd * gamma)
        return rnd(dx, self.dt)

    def bn2_conv2_bwd(self, loc_sum_dy, loc_sum_dy_xmu, sum_dy, sum_dy_xmu, n_total):
        p, dt = self.st.params, self.dt
        self.g["layer2.1.weight"] = loc_sum_dy_xmu * self.is2  # fp32, local (DDP averages it)
        self.g["layer2.1.bias"] = loc_sum_dy
        dy2 = self._bn_bwd_elemt(self.da2, self.y2, self.mu2, self.is2, p["layer2.1.weight"], sum_dy, sum_dy_xmu, n_total)
        w2 = rnd(p["layer2.0.weight"], dt)
        self.g["layer2.0.weight"] = rnd(torch.nn.grad.conv2d_weight(self.p1, w2.shape, dy2, 1, 2), dt)
        self.g["layer2.0.bias"] = rnd(dy2.sum(dim=(0, 2, 3)), dt)
        dp1 = rnd(torch.nn.grad.conv2d_input(self.p1.shape, w2, dy2, 1, 2), dt)
        self.da1 = _unpool_relu_bwd(dp1, self.arg1, self.a1)
        return self._bn_bwd_sums(self.da1, self.y1, self.mu1)

    def bn1_conv1_bwd(self, loc_sum_dy, loc_sum_dy_xmu, sum_dy, sum_dy_xmu, n_total):
        p, dt = self.st.params, self.dt
        self.g["layer1.1.weight"] = loc_sum_dy_xmu * self.is1
        self.g["layer1.1.bias"] = loc_sum_dy
        dy1 = self._bn_bwd_elemt(self.da1, self.y1, self.mu1, self.is1, p["layer1.1.weight"], sum_dy, sum_dy_xmu, n_total)
        self.g["layer1.0.weight"] = rnd(torch.nn.grad.conv2d_weight(self.x, (16, 1, 5, 5), dy1, 1, 2), dt)
        self.g["layer1.0.bias"] = rnd(dy1.sum(dim=(0, 2, 3)), dt)
        return self.g


# --------------------------------------------------------------------------- scaler / optimizer


def amp_update_scale(scale: float, tracker: int, found_inf: float, growth_factor=2.0, backoff_factor=0.5,
                     growth_interval=2000):
    """`_amp_update_scale_` (grad_scaler.py:549-556; semantics SURVEY A.3), fp32."""
    s = torch.tensor(scale, dtype=torch.float32)
    if found_inf >= 1.0:
        return float(s * torch.tensor(backoff_factor, dtype=torch.float32)), 0
    tracker += 1
    if tracker == growth_interval:
        grown = s * torch.tensor(growth_factor, dtype=torch.float32)
        if torch.isfinite(grown):
            s = grown
        tracker = 0
    return float(s), tracker


def inv_scale_of(scale: float) -> torch.Tensor:
    """grad_scaler.py:351-355: reciprocal in fp64, result stored fp32."""
    return torch.tensor(scale, dtype=torch.float32).double().reciprocal().float()


@dataclass
class StepResult:
    losses: List[float]  # per rank, value of the `dt` loss scalar
    found_inf: float
    scaled_grads: torch.Tensor  # flat, after the all-reduce, before unscale (tap T-b)
    grads: torch.Tensor  # flat, after unscale_ (tap T-c)
    logits: List[torch.Tensor]


def train_step(st: OracleState, xs: Sequence[torch.Tensor], ys: Sequence[torch.Tensor], dtype: str = "fp16",
               use_scaler: bool = True, sync_bn: bool = True, lr: float = 1e-4, momentum: float = 0.0,
               weight_decay: float = 0.0, dampening: float = 0.0, nesterov: bool = False,
               accumulate: bool = False) -> StepResult:
    """One `train()` iteration on W = len(xs) ranks that share `st` (replicas are
    identical, as DDP guarantees).  Mutates `st` like the reference mutates model,
    optimizer and scaler.  With W > 1, `sync_bn=True` is the reference behaviour
    (ddp_main.py:120); `sync_bn=False` gives plain per-rank BN with DDP's rank-0
    buffer broadcast."""
    dt = _DT[dtype]
    W = len(xs)
    ranks = [_RankPass(st, x, y, dt) for x, y in zip(xs, ys)]
    p = st.params

    def fwd_stats(local, layer):
        means, invstds, counts = zip(*local)
        if W > 1 and sync_bn:
            mean, invstd, m2, n = merge_stats(means, invstds, counts)
            use = [(mean, invstd)] * W
        else:
            use = [(m, i) for m, i in zip(means, invstds)]
            # rank 0's statistics define the running buffers every rank ends up with
            mean, n = means[0], float(counts[0])
            m2 = (1.0 / (invstds[0] * invstds[0]) - BN_EPS) * n
        st.running_mean[layer] = (1 - BN_MOMENTUM) * st.running_mean[layer] + BN_MOMENTUM * mean
        st.running_var[layer] = (1 - BN_MOMENTUM) * st.running_var[layer] + BN_MOMENTUM * (m2 / (n - 1))
        st.num_batches_tracked[layer] += 1
        return use

    use1 = fwd_stats([r.conv1() for r in ranks], 0)
    use2 = fwd_stats([r.bn1_conv2(*u) for r, u in zip(ranks, use1)], 1)
    losses = [float(r.bn2_head(*u)) for r, u in zip(ranks, use2)]

    scale = st.scale if use_scaler else 1.0

    def bwd_sums(local, counts):
        if W > 1 and sync_bn:
            # all_reduce(SUM) of the two vectors (_functions.py:159), fixed rank order 0..W-1
            s0, s1 = local[0][0].clone(), local[0][1].clone()
            for l in local[1:]:
                s0, s1 = s0 + l[0], s1 + l[1]
            return [(l[0], l[1], s0, s1, float(sum(counts))) for l in local]
        return [(l[0], l[1], l[0], l[1], float(c)) for l, c in zip(local, counts)]

    n2 = [r.B * 14 * 14 for r in ranks]
    n1 = [r.B * 28 * 28 for r in ranks]
    b2 = bwd_sums([r.head_bwd(scale) for r in ranks], n2)
    b1 = bwd_sums([r.bn2_conv2_bwd(*a) for r, a in zip(ranks, b2)], n1)
    per_rank = [flatten_grads(r.bn1_conv1_bwd(*a)) for r, a in zip(ranks, b1)]
    # gradient accumulation: autograd ADDS this backward's (rounded, scaled) gradient to what `.grad` already holds, in
    # fp32.  `accumulate=True` is a backward inside `with model.no_sync():` -- nothing is reduced, unscaled or applied,
    # BN statistics / running buffers above were updated as in any forward (distributed.py:1441-1470)
    if st.grad_accum is not None:
        per_rank = [a + g for a, g in zip(st.grad_accum, per_rank)]
    if accumulate:
        st.grad_accum = per_rank
        return StepResult(losses, float("nan"), per_rank[0].clone(), per_rank[0].clone(), [r.logits for r in ranks])
    st.grad_accum = None

    # DDP Reducer: bucket = grad * (1/W), then all-reduce SUM (rank order here)
    inv_w = torch.tensor(1.0 / W, dtype=torch.float32)
    avg = per_rank[0] * inv_w
    for g in per_rank[1:]:
        avg = avg + g * inv_w
    scaled = avg.clone()

    if use_scaler:
        avg = avg * inv_scale_of(st.scale)  # every element, even when non-finite ones exist
        found_inf = 0.0 if bool(torch.isfinite(avg).all()) else 1.0
    else:
        found_inf = 0.0
    if found_inf == 0.0:
        upd = avg
        if momentum != 0.0 or weight_decay != 0.0:
            # the other SGD knobs, op for op as `_single_tensor_sgd` / `_multi_tensor_sgd` (sgd.py:330-465); a step the
            # scaler skips touches neither the parameters nor the buffer, and `.grad` (tap T-c) stays as it was
            flat = st.flat_params()
            upd = avg.add(flat, alpha=weight_decay) if weight_decay != 0.0 else avg.clone()
            if momentum != 0.0:
                if st.momentum_buffer is None:
                    st.momentum_buffer = upd.clone()
                else:
                    st.momentum_buffer.mul_(momentum).add_(upd, alpha=1.0 - dampening)
                upd = upd.add(st.momentum_buffer, alpha=momentum) if nesterov else st.momentum_buffer
        off = 0
        for name, shape in PARAM_SPECS:
            n = math.prod(shape)
            p[name] = p[name] + torch.tensor(-lr, dtype=torch.float32) * upd[off:off + n].view(shape)
            off += n
    if use_scaler:
        st.scale, st.growth_tracker = amp_update_scale(st.scale, st.growth_tracker, found_inf, st.growth_factor,
                                                       st.backoff_factor, st.growth_interval)
    return StepResult(losses, found_inf, scaled, avg, [r.logits for r in ranks])


# --------------------------------------------------------------------------- eval / test()


def eval_logits(st: OracleState, x: torch.Tensor, dtype: str = "fp16") -> torch.Tensor:
    """`model.eval()` forward (ddp_main.py:98-105): BN uses the running buffers."""
    dt = _DT[dtype]
    p = st.params
    h = rnd(x.float(), dt)
    for li, (cw, cb, g, b) in enumerate([("layer1.0.weight", "layer1.0.bias", "layer1.1.weight", "layer1.1.bias"),
                                          ("layer2.0.weight", "layer2.0.bias", "layer2.1.weight", "layer2.1.bias")]):
        yv = _conv(h, p[cw], p[cb], dt)
        invstd = 1.0 / torch.sqrt(st.running_var[li] + BN_EPS)
        a = _bn_apply(yv, st.running_mean[li], invstd, p[g], p[b], dt)
        h, _ = _relu_pool(a)
    flat = h.reshape(x.shape[0], -1)
    return rnd(flat @ rnd(p["fc.weight"], dt).t() + rnd(p["fc.bias"], dt), dt)


def count_correct(logits: torch.Tensor, labels: torch.Tensor):
    """ddp_main.py:106-107: `size += B`, `correct += (argmax(1) == labels).float().sum()`;
    argmax takes the first maximum and treats NaN as the maximum."""
    return float((logits.argmax(1) == labels).to(torch.float32).sum()), float(logits.shape[0])


def test_counts(st: OracleState, shards: Sequence[Sequence], dtype: str = "fp16"):
    """`test()` on W ranks: per-rank fp32 accumulators, then reduce(SUM) to rank 0
    (ddp_main.py:108-109).  `shards[r]` is that rank's list of (images, labels)."""
    tot_c = tot_s = 0.0
    for batches in shards:
        c = s = 0.0
        for x, y in batches:
            dc, ds = count_correct(eval_logits(st, x, dtype), y)
            c += dc
            s += ds
        tot_c += c
        tot_s += s
    return tot_c, tot_s


# --------------------------------------------------------------------------- synthetic data


def synthetic_mnist(n: int, seed: int = 3407):
    """MNIST-shaped data with ToTensor's quantisation: uint8/255 in [0,1]."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n, 1, 28, 28), dtype=torch.uint8, generator=g).float() / 255
    y = torch.randint(0, 10, (n,), generator=g)
    return x, y
