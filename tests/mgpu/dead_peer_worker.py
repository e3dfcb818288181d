"""torchrun worker (world 2): a peer that stops stepping must end in DDPSTEP_E_PEER on the waiting rank -- within the
configured timeout, with no update applied by the voided step -- never in a hang or in a silently different replica.
The reference's NCCL would block until its watchdog fires."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from ddp_practice_b200 import Engine, NativeError  # noqa: E402
from oracle import convnet_oracle as O  # noqa: E402
from tests.refmodel import RefConvNet, bn_buffers, flat_params  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.manual_seed(7)
    model = RefConvNet(torch.float16, "cuda")
    eng = Engine(dtype="fp16", batch_cap=32, rank=rank, world=world, device=local)
    eng.connect()
    eng.load_state(flat_params(model).numpy(), bn_buffers(model).numpy())
    x, y = O.synthetic_mnist(32 * 4 * world)
    x, y = x[rank::world].cuda(), y[rank::world].cuda()
    for s in range(2):  # healthy steps
        eng.train_step(x[s * 32:(s + 1) * 32].contiguous(), y[s * 32:(s + 1) * 32].contiguous())
    torch.cuda.synchronize()
    eng.check_device()
    eng.set_peer_timeout_ms(300.0)  # the healthy steps ran under the default (20 s): graph capture skews the first one
    dist.barrier()
    out = {"ok": True}
    if rank == 0:  # rank 1 "dies": it never enters step 2
        before = torch.cat([eng.params, eng.bn_buffers, eng.scale, eng.growth_tracker.float()]).clone()
        t0 = time.perf_counter()
        eng.train_step(x[64:96].contiguous(), y[64:96].contiguous())
        torch.cuda.synchronize()
        waited = time.perf_counter() - t0
        assert waited < 5.0, f"the voided step took {waited:.2f} s"
        try:
            eng.check_device()
            raise AssertionError("ddpstep_device_status did not report the dead peer")
        except NativeError as e:
            assert "-4" in str(e) and "peer" in str(e), str(e)
        after = torch.cat([eng.params, eng.bn_buffers, eng.scale, eng.growth_tracker.float()])
        assert torch.equal(before.view(torch.int32), after.view(torch.int32)), "the voided step changed state"
        try:  # the step entry points refuse from now on, without a device synchronisation
            eng.train_step(x[96:128].contiguous(), y[96:128].contiguous())
            raise AssertionError("a step was accepted after a peer timeout")
        except NativeError as e:
            assert "-4" in str(e), str(e)
        out["waited_s"] = waited
    dist.barrier()
    if rank == 0:
        print(json.dumps(out), flush=True)
    eng.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
