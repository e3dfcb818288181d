"""torchrun worker: ddp_practice_b200.comm_hook on an arbitrary model under torch's own DDP (SURVEY 8b alternative B,
8f N3): bucket_cap_mb forces several buckets; gradients must equal stock DDP's (NCCL) and be bit-identical on every rank."""
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from ddp_practice_b200.comm_hook import B200AllreduceState, b200_allreduce_hook  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.backends.cuda.matmul.allow_tf32 = False

    def make():
        torch.manual_seed(7)
        return nn.Sequential(nn.Linear(64, 600), nn.ReLU(), nn.Linear(600, 600), nn.ReLU(), nn.Linear(600, 10)).cuda()

    plain = nn.parallel.DistributedDataParallel(make(), device_ids=[local], bucket_cap_mb=0.2)
    hooked = nn.parallel.DistributedDataParallel(make(), device_ids=[local], bucket_cap_mb=0.2)
    state = B200AllreduceState.create(device=local)
    hooked.register_comm_hook(state, b200_allreduce_hook)
    torch.manual_seed(100 + rank)
    worst = 0.0
    for it in range(3):
        x, y = torch.randn(16, 64, device="cuda"), torch.randint(0, 10, (16,), device="cuda")
        for m in (plain, hooked):
            m.zero_grad()
            nn.functional.cross_entropy(m(x), y).backward()
        torch.cuda.synchronize()
        state.engine.check_device()
        gp = torch.cat([p.grad.reshape(-1) for p in plain.parameters()])
        gh = torch.cat([p.grad.reshape(-1) for p in hooked.parameters()])
        worst = max(worst, ((gp - gh).norm() / gp.norm()).item())
        assert torch.allclose(gp, gh, rtol=1e-5, atol=1e-8), (rank, it)
        other = gh.clone()
        dist.broadcast(other, 0)
        assert torch.equal(gh.view(torch.int32), other.view(torch.int32)), f"rank {rank}: hooked gradients differ from rank 0"
    assert state.buckets_seen >= 5  # 1 on the first iteration, several afterwards (Reducer::rebuild_buckets)
    if rank == 0:
        print(json.dumps({"ok": True, "world": world, "buckets_seen": state.buckets_seen, "worst_rel_l2_vs_nccl": worst}), flush=True)
    state.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
