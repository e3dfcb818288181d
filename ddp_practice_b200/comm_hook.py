"""The one-shot NVLink all-reduce as a DDP communication hook (SURVEY.md 8b alternative B, 8f N3).

`torch.nn.parallel.DistributedDataParallel.register_comm_hook(state, hook)` is the plugin point the reference's stack
already has (torch/nn/parallel/distributed.py:1987-2067): DDP hands every gradient bucket -- one or many, of ANY model,
whatever `bucket_cap_mb` and the parameter list make of it -- to `hook(state, bucket)` and expects a
`Future[Tensor]` holding the reduced bucket.  With a hook registered DDP does not pre-divide (distributed.py:2021-2023),
so the hook averages, like `default_hooks.allreduce_hook`.

    model = DistributedDataParallel(any_model.cuda(), device_ids=[local_rank])
    hook_state = B200AllreduceState.create()            # one native context per process: peer mapping over NVLink
    model.register_comm_hook(hook_state, b200_allreduce_hook)

This does NOT give the fused step (forward/backward stay torch's); it replaces exactly the ncclAllReduce behind
ddp_main.py:121-123 by the library's exchange (fixed rank-order sum: bit-identical replicas, deterministic)."""
from typing import Callable, Optional

import torch
import torch.distributed as dist


class B200AllreduceState:
    """`state` argument of the hook: who reduces a bucket.  `reduce_fn(tensor)` must average `tensor` over the ranks in
    place; by default that is Engine.allreduce_bucket (libddpstep.so, NVLink peer memory)."""

    def __init__(self, reduce_fn: Callable[[torch.Tensor], None], engine=None):
        self.reduce_fn, self.engine = reduce_fn, engine
        self.buckets_seen = 0

    @classmethod
    def create(cls, group=None, device: Optional[int] = None) -> "B200AllreduceState":
        from .engine import Engine
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        eng = Engine(dtype="fp32", batch_cap=1, rank=rank, world=world, device=device, use_scaler=False)
        eng.connect(group)
        return cls(lambda t: eng.allreduce_bucket(t, average=True), eng)

    def close(self):
        if self.engine is not None:
            self.engine.close()
            self.engine = None


def b200_allreduce_hook(state: B200AllreduceState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    """hook(state, bucket: dist.GradBucket) -> Future[Tensor], the contract of distributed.py:1987-2067."""
    buf = bucket.buffer()
    if buf.dtype != torch.float32 or not buf.is_contiguous():
        raise NotImplementedError("b200_allreduce_hook reduces contiguous fp32 buckets (DDP's default gradient dtype)")
    state.reduce_fn(buf)  # stream-ordered on the current stream: DDP consumes the future's value on that stream too
    state.buckets_seen += 1
    fut: torch.futures.Future = torch.futures.Future()
    fut.set_result(buf)
    return fut
