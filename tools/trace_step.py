"""Timeline of one training step from in-kernel timestamps (CTA 0, thread 0 of every kernel; build the tracing library
with ddp_practice_b200.build.build_tools()).  Single GPU:  python tools/trace_step.py [batch]
N GPUs:  torchrun --nproc-per-node N tools/trace_step.py [batch]   (rank 0 prints; SyncBN + DDP exchanges included)"""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
so = os.path.join(ROOT, "tools", "libddpstep_trace.so")
from ddp_practice_b200 import build, _lib
build.LIB_PATH = so; _lib.LIB_PATH = so
from ddp_practice_b200 import Engine
from oracle import convnet_oracle as O
from tests.refmodel import RefConvNet
lib = _lib.load(); lib.ddpstep_read_trace.argtypes = [C.c_void_p]
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.manual_seed(0)
st = O.OracleState.from_state_dict(RefConvNet(None, "cpu").state_dict())
e = Engine(dtype="fp16", batch_cap=B, rank=rank, world=world, device=local)
e.connect()
bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
e.load_state(st.flat_params().numpy(), bn.numpy())
x, y = O.synthetic_mnist(B * world); x, y = x[rank::world].cuda().contiguous(), y[rank::world].cuda().contiguous()
for i in range(50): e.train_step(x, y)
torch.cuda.synchronize()
e.check_device()
buf = (C.c_ulonglong * 256)(); lib.ddpstep_read_trace(buf); t = list(buf)
names = ["w2_tiles_refresh", "conv1_fwd", "bn1_relu_pool", "conv2_fwd", "bn2_relu_pool", "fc_ce", "fc_wgrad", "bn2_bwd_reduce",
         "conv2_wgrad", "bn2_bwd_apply", "conv2_dgrad", "bn1_bwd_reduce", "conv1_wgrad", "step", "conv2_wgrad_reduce", "fc_reduce_push"]
ev = [(t[16 * k], names[k], k) for k in range(len(names)) if t[16 * k]]
ev.sort(); base = ev[0][0]
if rank == 0:
    print(f"# world {world}, batch {B}/GPU, rank 0, last of 50 steps; [i]+x = in-kernel mark i at x us after the kernel's entry")
    for ts, n, k in ev:
        inner = " ".join(f"[{i}]+{(t[16*k+i]-ts)/1e3:.2f}" for i in range(1, 16) if t[16 * k + i])
        print(f"{n:18s} +{(ts-base)/1e3:7.2f} us   {inner}")
e.close()
if world > 1:
    dist.barrier(); dist.destroy_process_group()
