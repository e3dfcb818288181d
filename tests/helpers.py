"""Shared test helpers: golden loading, oracle state construction, error metrics."""
import math
import os

import numpy as np
import torch

from oracle.convnet_oracle import PARAM_SPECS, OracleState

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name)))


def state_from_golden(g, prefix="init_"):
    flat = torch.from_numpy(g[f"{prefix}params"]).clone()
    params, off = {}, 0
    for n, s in PARAM_SPECS:
        k = math.prod(s)
        params[n] = flat[off:off + k].view(s).clone()
        off += k
    return OracleState(
        params=params,
        running_mean=[torch.from_numpy(g[f"{prefix}rm{i}"]).clone() for i in range(2)],
        running_var=[torch.from_numpy(g[f"{prefix}rv{i}"]).clone() for i in range(2)],
        num_batches_tracked=[int(g[f"{prefix}nbt{i}"]) for i in range(2)],
    )


def param_slices():
    out, off = {}, 0
    for n, s in PARAM_SPECS:
        k = math.prod(s)
        out[n] = slice(off, off + k)
        off += k
    return out


# conv biases feed straight into a BatchNorm, so their gradient is mathematically
# zero and what both sides hold is pure rounding noise: compare them on the scale
# of the sibling weight gradient instead of their own.
_NOISE_ONLY = {"layer1.0.bias": "layer1.0.weight", "layer2.0.bias": "layer2.0.weight"}


def grad_errors(got, ref, metric="l2"):
    """Per-parameter relative error of a flat 29,034-element gradient.
    metric "l2":  ||got-ref||_2 / ||ref||_2      (the stated relative tolerance is on this)
    metric "max": max|got-ref| / max|ref|
    For the two noise-only conv biases the denominator comes from the sibling weight."""
    got = torch.as_tensor(np.asarray(got), dtype=torch.float64)
    ref = torch.as_tensor(np.asarray(ref), dtype=torch.float64)
    sl = param_slices()
    norm = (lambda t: t.norm().item()) if metric == "l2" else (lambda t: t.abs().max().item())
    errs = {}
    for n, s in sl.items():
        scale = norm(ref[sl[_NOISE_ONLY.get(n, n)]])
        errs[n] = norm(got[s] - ref[s]) / max(scale, 1e-30)
    return errs


def assert_grads_close(got, ref, tol, what="", metric="l2"):
    errs = grad_errors(got, ref, metric)
    bad = {k: v for k, v in errs.items() if not v <= tol}
    assert not bad, f"{what} gradient mismatch ({metric}, tol {tol}): {bad}"


# ---- discrete routing (ReLU mask + max-pool winner) --------------------------------------
# Two correct implementations that differ in the last ulp of a conv output can disagree on
# which element of a 2x2 window wins the pool (or whether a ReLU is open) when two values
# tie at the storage precision.  Each such flip reroutes one whole gradient path (~1/B of a
# layer's gradient signal), so gradient tolerances are stated for "routing agrees" and
# relaxed, with the flip count reported, when it does not.


def routing_of(a):
    """a: BN output [B,C,H,W] (fp32 values).  Returns (winner index 0..3, live mask)."""
    from oracle.convnet_oracle import _relu_pool
    p, arg = _relu_pool(a)
    return arg, p > 0


def bn_out_from_conv(y, gamma, beta, dtype):
    """BN(train) output recomputed from a conv output with the oracle's formulas."""
    from oracle import convnet_oracle as O
    mean, invstd, _ = O._local_stats(y)
    return O._bn_apply(y, mean, invstd, gamma, beta, O._DT[dtype])


def count_flips(r_a, r_b):
    (arg_a, live_a), (arg_b, live_b) = r_a, r_b
    return int(((live_a != live_b) | (live_a & live_b & (arg_a != arg_b))).sum())
