"""Host-side owner of one rank's native context (libddpstep.so, include/ddpstep.h).

torch is plumbing here: device selection, the current stream, zero-copy tensor views of
the library's device state, and `torch.distributed` for the one-time exchange of IPC
handles.  All arithmetic of the training step happens in the CUDA library."""
import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _lib
from ._lib import DTYPES, HANDLE_BYTES, N_BNBUF, N_PARAMS

TORCH_DT = {"fp32": torch.float32, "fp16": torch.float16, "bf16": torch.bfloat16}


class _DevArray:
    """Minimal __cuda_array_interface__ carrier for a device pointer owned by the library."""

    def __init__(self, ptr, shape, typestr, owner):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}
        self._owner = owner  # keep the engine alive as long as a view exists


class NativeError(RuntimeError):
    pass


class Engine:
    """One rank of the fused DDP+AMP ConvNet step.

    Mirrors the objects `ddp_main.py:117-126` builds (model replica, DDP reducer, SGD,
    GradScaler) as a single native context; `train_step` is one iteration of
    `ddp_main.py:86-93`, `eval_batch` one iteration of `ddp_main.py:101-107`."""

    def __init__(self, dtype: str = "fp16", batch_cap: int = 32, rank: int = 0, world: int = 1,
                 device: Optional[int] = None, use_scaler: Optional[bool] = None, sync_bn: bool = True,
                 use_graph: bool = True, lr: float = 1e-4, init_scale: float = 65536.0, growth_factor: float = 2.0,
                 backoff_factor: float = 0.5, growth_interval: int = 2000, conv_bias_double_round: bool = True,
                 momentum: float = 0.0, weight_decay: float = 0.0, dampening: float = 0.0, nesterov: bool = False):
        if not torch.cuda.is_available():
            raise NativeError("ddp_practice_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.dtype = dtype
        self.device = torch.cuda.current_device() if device is None else int(device)
        cfg = _lib.Config()
        self.lib.ddpstep_default_config(C.byref(cfg))
        cfg.rank, cfg.world, cfg.device = rank, world, self.device
        cfg.dtype, cfg.batch_cap = DTYPES[dtype], batch_cap
        cfg.use_scaler = int(dtype != "fp32") if use_scaler is None else int(use_scaler)
        cfg.sync_bn, cfg.use_graph = int(sync_bn), int(use_graph)
        cfg.conv_bias_double_round = int(conv_bias_double_round)
        cfg.lr, cfg.init_scale = lr, init_scale
        cfg.growth_factor, cfg.backoff_factor, cfg.growth_interval = growth_factor, backoff_factor, growth_interval
        cfg.momentum, cfg.weight_decay, cfg.dampening, cfg.nesterov = momentum, weight_decay, dampening, int(nesterov)
        self.cfg = cfg
        self.rank, self.world, self.batch_cap = rank, world, batch_cap
        ctx = C.c_void_p()
        rc = self.lib.ddpstep_create(C.byref(cfg), C.byref(ctx))
        if rc != 0:
            raise NativeError(f"ddpstep_create failed ({rc}): {self.lib.ddpstep_last_error(None).decode()}")
        self.ctx = ctx
        p = _lib.Ptrs()
        self._check(self.lib.ddpstep_state_ptrs(self.ctx, C.byref(p)))
        self._ptrs = p
        self._loss_out = C.c_float()
        self._loss_ref = C.byref(self._loss_out)
        dev = f"cuda:{self.device}"
        view = lambda ptr, shape, ts: torch.as_tensor(_DevArray(ptr, shape, ts, self), device=dev)
        # zero-copy views of the library's state
        self.params = view(p.params, (N_PARAMS,), "<f4")
        self.grads = view(p.grads, (N_PARAMS,), "<f4")
        self.grads_scaled = view(p.grads_scaled, (N_PARAMS,), "<f4")
        self.bn_buffers = view(p.bn_buffers, (N_BNBUF,), "<f4")
        self.num_batches_tracked = view(p.num_batches_tracked, (2,), "<i8")
        self.scale = view(p.scale, (1,), "<f4")
        self.growth_tracker = view(p.growth_tracker, (1,), "<i4")
        self.found_inf = view(p.found_inf, (1,), "<f4")
        self.loss = view(p.loss, (1,), "<f4")
        self.counts = view(p.counts, (2,), "<f4")
        self.input_stage = view(p.input_stage, (batch_cap, 1, 28, 28), "<f4")
        self.label_stage = view(p.label_stage, (batch_cap,), "<i8")
        self.momentum_buffer = view(p.momentum_buffer, (N_PARAMS,), "<f4")  # optimizer.state[p]['momentum_buffer'], flat
        self.momentum_ready = view(p.momentum_ready, (1,), "<i4")           # 1 once a step has created them
        if dtype == "fp32":
            self.logits = view(p.logits, (batch_cap, 10), "<f4")
        else:
            self.logits = view(p.logits, (batch_cap, 10), "<i2").view(TORCH_DT[dtype])

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc):
        if rc != 0:
            raise NativeError(f"libddpstep error {rc}: {self.lib.ddpstep_last_error(self.ctx).decode()}")

    def close(self):
        if getattr(self, "ctx", None):
            self.lib.ddpstep_destroy(self.ctx)
            self.ctx = None

    @staticmethod
    def _stream():
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def connect(self, group=None):
        """All-gather the IPC handles of every rank's symmetric region over the caller's
        control plane (`torch.distributed`), then map the peers.  Collective."""
        if self.world == 1:
            return
        import torch.distributed as dist
        buf = (C.c_ubyte * HANDLE_BYTES)()
        self._check(self.lib.ddpstep_export_handle(self.ctx, buf))
        gathered = [None] * self.world
        dist.all_gather_object(gathered, bytes(buf), group=group)
        blob = b"".join(gathered)
        assert len(blob) == HANDLE_BYTES * self.world
        self._check(self.lib.ddpstep_import_peers(self.ctx, blob, self.world))
        dist.barrier(group=group)

    # ------------------------------------------------------------------ state
    def load_state(self, params, bn_buffers, nbt=None):
        pa = np.ascontiguousarray(np.asarray(params, dtype=np.float32).reshape(-1))
        bn = np.ascontiguousarray(np.asarray(bn_buffers, dtype=np.float32).reshape(-1))
        assert pa.size == N_PARAMS and bn.size == N_BNBUF
        nb = None if nbt is None else np.ascontiguousarray(np.asarray(nbt, dtype=np.int64))
        self._check(self.lib.ddpstep_load_state(self.ctx, pa.ctypes.data, bn.ctypes.data,
                                                None if nb is None else nb.ctypes.data))

    def read_state(self):
        pa, bn = np.empty(N_PARAMS, np.float32), np.empty(N_BNBUF, np.float32)
        nb = np.empty(2, np.int64)
        sc, tr = C.c_float(), C.c_int32()
        self._check(self.lib.ddpstep_read_state(self.ctx, pa.ctypes.data, bn.ctypes.data, nb.ctypes.data,
                                                C.byref(sc), C.byref(tr)))
        return {"params": pa, "bn_buffers": bn, "nbt": nb, "scale": sc.value, "growth_tracker": tr.value}

    def set_scaler(self, scale: float, growth_tracker: int = 0):
        self._check(self.lib.ddpstep_set_scaler(self.ctx, scale, growth_tracker))

    # ------------------------------------------------------------------ the hot path
    def train_step(self, x: torch.Tensor, y: torch.Tensor):
        """x: cuda fp32 [B,1,28,28] contiguous, y: cuda int64 [B].  Asynchronous."""
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
        assert y.is_cuda and y.dtype == torch.int64 and y.is_contiguous()
        self._check(self.lib.ddpstep_train_step(self.ctx, x.data_ptr(), y.data_ptr(), x.shape[0], self._stream()))

    def train_steps(self, x: torch.Tensor, y: torch.Tensor, batch: int):
        """x.shape[0] // batch consecutive training steps on device-resident batches as ONE graph launch."""
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous() and x.shape[0] % batch == 0
        assert y.is_cuda and y.dtype == torch.int64 and y.is_contiguous() and y.shape[0] == x.shape[0]
        self._check(self.lib.ddpstep_train_steps(self.ctx, x.data_ptr(), y.data_ptr(), x.shape[0] // batch, batch, self._stream()))

    def train_steps_indexed(self, images_u8: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor, batch: int):
        """An epoch segment: idx holds the sampler's indices of idx.shape[0] // batch full batches; one graph launch."""
        assert images_u8.is_cuda and images_u8.dtype == torch.uint8 and images_u8.is_contiguous()
        assert labels.is_cuda and labels.dtype == torch.int64 and idx.is_cuda and idx.dtype == torch.int64
        assert idx.is_contiguous() and idx.shape[0] % batch == 0
        self._check(self.lib.ddpstep_train_steps_indexed(self.ctx, images_u8.data_ptr(), labels.data_ptr(), idx.data_ptr(),
                                                         idx.shape[0] // batch, batch, self._stream()))

    def train_step_accumulate(self, x: torch.Tensor, y: torch.Tensor):
        """A micro-step inside `with model.no_sync():` -- forward + backward, the gradient accumulated locally."""
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
        assert y.is_cuda and y.dtype == torch.int64 and y.is_contiguous()
        self._check(self.lib.ddpstep_train_step_accumulate(self.ctx, x.data_ptr(), y.data_ptr(), x.shape[0], self._stream()))

    def train_step_host(self, x: torch.Tensor, y: torch.Tensor, want_loss: bool = True):
        """Same from (pinned) host tensors; returns the loss (synchronises) if asked."""
        if x.is_cuda or y.is_cuda or x.dtype != torch.float32 or y.dtype != torch.int64 or not (
                x.is_contiguous() and y.is_contiguous()):
            raise ValueError("train_step_host wants contiguous CPU tensors: fp32 images, int64 labels")
        loss = self._loss_out
        rc = self.lib.ddpstep_train_step_host(self.ctx, x.data_ptr(), y.data_ptr(), x.shape[0],
                                              self._loss_ref if want_loss else None, self._stream())
        if rc != 0:
            self._check(rc)
        return loss.value if want_loss else None

    def train_step_indexed(self, images_u8: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor):
        """Input-pipeline form: resident uint8 dataset [N,1,28,28] / [N,784], labels [N] int64, and this batch's
        sampler indices [B] int64 (all on the GPU).  Gather and ToTensor's /255 are fused into conv1's load."""
        assert images_u8.is_cuda and images_u8.dtype == torch.uint8 and images_u8.is_contiguous()
        assert labels.is_cuda and labels.dtype == torch.int64 and idx.is_cuda and idx.dtype == torch.int64
        assert idx.is_contiguous() and labels.is_contiguous()
        self._check(self.lib.ddpstep_train_step_indexed(self.ctx, images_u8.data_ptr(), labels.data_ptr(), idx.data_ptr(),
                                                        idx.shape[0], self._stream()))

    def eval_batch_indexed(self, images_u8: torch.Tensor, labels: Optional[torch.Tensor], idx: torch.Tensor) -> torch.Tensor:
        assert images_u8.is_cuda and images_u8.dtype == torch.uint8 and idx.is_cuda and idx.dtype == torch.int64
        self._check(self.lib.ddpstep_eval_batch_indexed(self.ctx, images_u8.data_ptr(),
                                                        None if labels is None else labels.data_ptr(), idx.data_ptr(),
                                                        idx.shape[0], self._stream()))
        return self.logits[: idx.shape[0]]

    def eval_batch(self, x: torch.Tensor, y: Optional[torch.Tensor] = None) -> torch.Tensor:
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
        if y is not None:
            assert y.is_cuda and y.dtype == torch.int64 and y.is_contiguous()
        self._check(self.lib.ddpstep_eval_batch(self.ctx, x.data_ptr(), None if y is None else y.data_ptr(),
                                                x.shape[0], self._stream()))
        return self.logits[: x.shape[0]]

    def reset_counts(self):
        self._check(self.lib.ddpstep_reset_counts(self.ctx, self._stream()))

    def reduce_counts(self):
        self._check(self.lib.ddpstep_reduce_counts(self.ctx, self._stream()))

    def read_counts(self):
        c, s = C.c_float(), C.c_float()
        self._check(self.lib.ddpstep_read_counts(self.ctx, C.byref(c), C.byref(s), self._stream()))
        return c.value, s.value

    # ------------------------------------------------------------------ taps
    ACT = {"y1": (0, (16, 28, 28)), "p1": (1, (16, 14, 14)), "y2": (2, (32, 14, 14)), "p2": (3, (32, 7, 7)),
           "dlogits": (4, (10,)), "dy2": (5, (32, 14, 14)), "dp1": (6, (16, 14, 14))}

    def read_activation(self, name: str, batch: int) -> torch.Tensor:
        which, shape = self.ACT[name]
        tdt = TORCH_DT[self.dtype]
        out = torch.empty((batch, *shape), dtype=tdt)
        self._check(self.lib.ddpstep_read_activation(self.ctx, which, out.data_ptr(), out.numel() * out.element_size()))
        return out.float()

    def profile_step(self, x: torch.Tensor, y: torch.Tensor, in_graph: bool = True):
        """One training step with per-kernel device timing: {kernel name: ms}.  A negative value = that kernel did not
        run at this batch (its work is fused into another one); its magnitude is the bare event-to-event interval."""
        ms = (C.c_float * 16)()
        self._check(self.lib.ddpstep_profile_step(self.ctx, x.data_ptr(), y.data_ptr(), x.shape[0], int(in_graph),
                                                  ms, 16, self._stream()))
        out, i = {}, 0
        while self.lib.ddpstep_kernel_name(i):
            out[self.lib.ddpstep_kernel_name(i).decode()] = float(ms[i])
            i += 1
        return out

    def set_sgd(self, lr: float, momentum: float = 0.0, weight_decay: float = 0.0, dampening: float = 0.0,
                nesterov: bool = False):
        """`optimizer.param_groups[0]` changed (LR scheduler): new values for the steps enqueued from now on."""
        self._check(self.lib.ddpstep_set_sgd(self.ctx, lr, momentum, weight_decay, dampening, int(nesterov), self._stream()))

    def set_sgd_group(self, tensor_mask: int, lr: float, momentum: float = 0.0, weight_decay: float = 0.0,
                      dampening: float = 0.0, nesterov: bool = False):
        """One of several `optimizer.param_groups`: bit i of tensor_mask = parameter tensor i in definition order."""
        self._check(self.lib.ddpstep_set_sgd_group(self.ctx, tensor_mask, lr, momentum, weight_decay, dampening,
                                                   int(nesterov), self._stream()))

    def set_peer_timeout_ms(self, ms: float):
        self._check(self.lib.ddpstep_set_peer_timeout_ms(self.ctx, float(ms)))

    def allreduce_only(self):
        """The tail kernel alone (all-reduce + unscale + SGD + scale update) on the last step's partial gradients."""
        self._check(self.lib.ddpstep_allreduce_only(self.ctx, self._stream()))

    def allreduce_bucket(self, bucket: torch.Tensor, average: bool = True):
        """In-place mean (or sum) over ranks of any contiguous fp32 CUDA tensor: a DDP comm hook's work for one bucket."""
        assert bucket.is_cuda and bucket.dtype == torch.float32 and bucket.is_contiguous()
        self._check(self.lib.ddpstep_allreduce_bucket(self.ctx, bucket.data_ptr(), bucket.numel(), int(average), self._stream()))

    def launch_count(self) -> int:
        return int(self.lib.ddpstep_launch_count(self.ctx))

    def check_device(self):
        self._check(self.lib.ddpstep_device_status(self.ctx))
