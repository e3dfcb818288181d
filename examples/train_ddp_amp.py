#!/usr/bin/env python
"""End-to-end example: DDP + AMP training of the DDP-practice ConvNet on the fused B200 step.

Same call surface as the reference's ddp_main_torchrun.py (DDP(ConvNet().cuda(), device_ids=[rank]),
autocast() in forward, scaler.scale(loss).backward(), scaler.step, scaler.update, DistributedSampler,
dist.reduce in the evaluation) -- only the two imports marked below differ.  MNIST cannot be
downloaded on the GPU box, so the data are synthetic MNIST-shaped tensors.

    python examples/train_ddp_amp.py                                   # 1 GPU
    torchrun --standalone --nproc-per-node 2 examples/train_ddp_amp.py  # 2 GPUs
"""
import argparse
import os
import sys
import time

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader, TensorDataset
from torch.utils.data.distributed import DistributedSampler

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ddp_practice_b200.shim import nn            # <- instead of `import torch.nn as nn`
from ddp_practice_b200.shim import GradScaler    # <- instead of `from torch.cuda.amp import GradScaler`


class ConvNet(nn.Module):
    def __init__(self, classes=10):
        super().__init__()
        stage = lambda cin, cout: nn.Sequential(nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2),
                                                nn.BatchNorm2d(cout), nn.ReLU(), nn.MaxPool2d(kernel_size=2, stride=2))
        self.layer1, self.layer2 = stage(1, 16), stage(16, 32)
        self.fc = nn.Linear(7 * 7 * 32, classes)

    def forward(self, x):
        with torch.cuda.amp.autocast():
            h = self.layer2(self.layer1(x))
            return self.fc(h.reshape(h.size(0), -1))


def synthetic(n, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (n, 1, 28, 28), dtype=torch.uint8, generator=g).float() / 255
    return TensorDataset(x, torch.randint(0, 10, (n,), generator=g))


def fit_one_epoch(model, loader, criterion, optimizer, scaler):
    model.train()
    for images, labels in loader:
        images, labels = images.cuda(), labels.cuda()
        outputs = model(images)
        loss = criterion(outputs, labels)
        optimizer.zero_grad()
        scaler.scale(loss).backward()
        scaler.step(optimizer)
        scaler.update()


def evaluate(model, loader, rank):
    model.eval()
    size, correct = torch.tensor(0.0).cuda(), torch.tensor(0.0).cuda()
    for images, labels in loader:
        images, labels = images.cuda(), labels.cuda()
        with torch.no_grad():
            outputs = model(images)
            size += images.size(0)
        correct += (outputs.argmax(1) == labels).type(torch.float).sum()
    if dist.is_initialized():
        dist.reduce(size, 0, op=dist.ReduceOp.SUM)
        dist.reduce(correct, 0, op=dist.ReduceOp.SUM)
    if rank == 0:
        print(f"Accuracy is {correct / size:.2%} ({int(correct)}/{int(size)})")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-e", "--epochs", type=int, default=1)
    ap.add_argument("-b", "--batch_size", type=int, default=32)
    ap.add_argument("--train-size", type=int, default=6400)
    args = ap.parse_args()
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    distributed = "RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1
    torch.cuda.set_device(local_rank)
    if distributed:
        dist.init_process_group(backend="nccl")
    model = ConvNet().cuda()
    model = nn.SyncBatchNorm.convert_sync_batchnorm(model)
    model = nn.parallel.DistributedDataParallel(model, device_ids=[local_rank])
    criterion = nn.CrossEntropyLoss().cuda()
    optimizer = torch.optim.SGD(model.parameters(), 1e-4)
    scaler = GradScaler()
    train, test = synthetic(args.train_size, 3407), synthetic(1000, 7)
    ts = DistributedSampler(train) if distributed else None
    es = DistributedSampler(test) if distributed else None
    tl = DataLoader(train, batch_size=args.batch_size, shuffle=False, sampler=ts, pin_memory=True)
    el = DataLoader(test, batch_size=args.batch_size, shuffle=False, sampler=es, pin_memory=True)
    t0 = time.time()
    for epoch in range(args.epochs):
        if local_rank == 0:
            print(f"begin training of epoch {epoch + 1}/{args.epochs}")
        if ts is not None:
            ts.set_epoch(epoch)
        fit_one_epoch(model, tl, criterion, optimizer, scaler)
    evaluate(model, el, local_rank)
    if local_rank == 0:
        torch.save({"model": model.state_dict(), "scaler": scaler.state_dict()}, "/tmp/ddp_b200_checkpoint.pt")
        print(f"scaler: {scaler.state_dict()}  time {time.time() - t0:.2f}s")
    if distributed:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
