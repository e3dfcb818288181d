"""The reference's own CPU path, restated so it can travel to the GPU box.  TEST / BASELINE
INFRASTRUCTURE ONLY (see oracle/convnet_oracle.py for who may import oracle/).

BASELINE.json configs[0]: "origin_main.py ConvNet ... batch 32, fp32 SGD lr 1e-4, on CPU
(reference path, .cuda() mapped to cpu)".  The reference is three Python scripts whose
arithmetic is torch's, so its CPU implementation is exactly: the layer stack of
origin_main.py:9-31 built from torch.nn, and the loop body of origin_main.py:60-66, run by
ATen's CPU kernels on the host cores.  /root/reference itself does not exist on the GPU box;
tests/test_oracle_golden.py pins this restatement to the real thing through the golden
vectors (same step, same numbers)."""
import time

import torch
import torch.nn as nn


def build_model(seed: int = 3407) -> nn.Module:
    torch.manual_seed(seed)

    def block(cin, cout):
        return nn.Sequential(nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2), nn.BatchNorm2d(cout),
                             nn.ReLU(), nn.MaxPool2d(kernel_size=2, stride=2))

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.layer1, self.layer2 = block(1, 16), block(16, 32)
            self.fc = nn.Linear(7 * 7 * 32, 10)

        def forward(self, x):
            h = self.layer2(self.layer1(x))
            return self.fc(h.reshape(h.size(0), -1))

    return Net()


class CpuReference:
    """model / criterion / optimizer of origin_main.py:85-87 and one call per loop iteration."""

    def __init__(self, threads: int = 0):
        if threads > 0:
            torch.set_num_threads(threads)
        self.threads = torch.get_num_threads()
        self.model = build_model()
        self.criterion = nn.CrossEntropyLoss()
        self.optimizer = torch.optim.SGD(self.model.parameters(), 1e-4)
        self.model.train()

    def step(self, images: torch.Tensor, labels: torch.Tensor) -> float:
        outputs = self.model(images)
        loss = self.criterion(outputs, labels)
        self.optimizer.zero_grad()
        loss.backward()
        self.optimizer.step()
        return loss

    def timed(self, x: torch.Tensor, y: torch.Tensor, batch: int, steps: int, warmup: int, budget_s: float = 1e9):
        """Run warmup + up to `steps` steps (stopping early after budget_s); seconds per step list."""
        nb = x.shape[0] // batch
        for i in range(warmup):
            j = i % nb
            self.step(x[j * batch:(j + 1) * batch], y[j * batch:(j + 1) * batch])
        times, t_all = [], time.perf_counter()
        for i in range(steps):
            j = i % nb
            t0 = time.perf_counter()
            loss = self.step(x[j * batch:(j + 1) * batch], y[j * batch:(j + 1) * batch])
            times.append(time.perf_counter() - t0)
            if time.perf_counter() - t_all > budget_s:
                break
        return times, float(loss.detach())
