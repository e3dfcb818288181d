"""The fused step as PyTorch operators: `torch.ops.ddp_b200.*` (visible to the dispatcher, usable from TorchScript-free
Python and from C++ through the dispatcher).  north_star words the delivery as "a PyTorch C++/CUDA extension behind a thin
C-ABI": the C ABI (include/ddpstep.h, libddpstep.so) is the product and carries no torch types; this module is the thin
adapter from tensors to its pointers, registered with `torch.library` instead of a compiled pybind module so that the
library itself stays free of libtorch.  An engine is addressed by an integer handle (`register(engine)`).

    h = torch_ops.register(Engine(dtype="fp16", batch_cap=32))
    loss = torch.ops.ddp_b200.train_step(images, labels, h)        # ddp_main.py:86-93, returns the loss tensor (T-a)
    logits = torch.ops.ddp_b200.eval_batch(images, h)              # ddp_main.py:104
    torch.ops.ddp_b200.allreduce_bucket_(bucket, h)                # the comm-hook exchange, in place
"""
import torch

from .engine import Engine

_ENGINES = {}
_lib = torch.library.Library("ddp_b200", "DEF")
_lib.define("train_step(Tensor images, Tensor labels, int handle) -> Tensor")
_lib.define("train_step_accumulate(Tensor images, Tensor labels, int handle) -> Tensor")
_lib.define("eval_batch(Tensor images, int handle) -> Tensor")
_lib.define("allreduce_bucket_(Tensor(a!) bucket, int handle) -> Tensor(a!)")


def register(engine: Engine) -> int:
    h = id(engine)
    _ENGINES[h] = engine
    return h


def unregister(handle: int) -> None:
    _ENGINES.pop(handle, None)


def _train_step(images, labels, handle):
    e = _ENGINES[handle]
    e.train_step(images.contiguous(), labels.contiguous())
    return e.loss  # zero-copy view of the library's loss word; valid once the stream has passed the step


def _train_step_accumulate(images, labels, handle):
    e = _ENGINES[handle]
    e.train_step_accumulate(images.contiguous(), labels.contiguous())
    return e.loss


def _eval_batch(images, handle):
    return _ENGINES[handle].eval_batch(images.contiguous())


def _allreduce_bucket_(bucket, handle):
    _ENGINES[handle].allreduce_bucket(bucket)
    return bucket


_lib.impl("train_step", _train_step, "CUDA")
_lib.impl("train_step_accumulate", _train_step_accumulate, "CUDA")
_lib.impl("eval_batch", _eval_batch, "CUDA")
_lib.impl("allreduce_bucket_", _allreduce_bucket_, "CUDA")
