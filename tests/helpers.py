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
