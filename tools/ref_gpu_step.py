"""The reference's own GPU path (torch eager: cuDNN/cuBLASLt/ATen + autocast + GradScaler + SGD; DDP + SyncBN over
NCCL when launched with torchrun), device-timed the same way bench.py times ours.  Context number only: the
reference arm of bench.py is the CPU path, per the tier rules."""
import json, os, statistics, sys
import torch, torch.nn as nn, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.refmodel import RefConvNet
from oracle.convnet_oracle import synthetic_mnist

def main():
    dtype = {"fp16": torch.float16, "bf16": torch.bfloat16}[sys.argv[1] if len(sys.argv) > 1 else "fp16"]
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 300
    world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1: dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.manual_seed(3407)
    model = RefConvNet(dtype, "cuda").cuda()
    if world > 1:
        model = nn.parallel.DistributedDataParallel(nn.SyncBatchNorm.convert_sync_batchnorm(model), device_ids=[local])
    criterion = nn.CrossEntropyLoss().cuda(); optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    scaler = torch.amp.GradScaler("cuda")
    x, y = synthetic_mnist(B * 16); x, y = x.cuda(), y.cuda()
    def step(i):
        j = i % 16
        images, labels = x[j*B:(j+1)*B], y[j*B:(j+1)*B]
        outputs = model(images); loss = criterion(outputs, labels)
        optimizer.zero_grad(); scaler.scale(loss).backward(); scaler.step(optimizer); scaler.update()
    model.train()
    for i in range(30): step(i)
    torch.cuda.synchronize()
    if world > 1: dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps): step(i)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = t.item()
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "torch-eager reference stack on GPU", "dtype": str(dtype), "batch_per_gpu": B, "n_gpus": world,
                          "ms_per_step": ms / steps, "img_per_s": world * B * steps / (ms * 1e-3)}), flush=True)
    if world > 1: dist.destroy_process_group()
main()
