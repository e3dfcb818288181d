"""The reference's model and loop bodies, rebuilt from torch.nn for tests that run where
/root/reference is not mounted (the GPU box).  Same layer stack and state-dict keys as
`ConvNet` in ddp_main.py:13-36 / origin_main.py:9-31; `amp_dtype=None` is the fp32 model
of origin_main.py, otherwise forward runs under autocast like ddp_main.py:31."""
import contextlib

import torch
import torch.nn as nn

from oracle.convnet_oracle import PARAM_SPECS


def _block(cin, cout):
    return nn.Sequential(nn.Conv2d(cin, cout, kernel_size=5, stride=1, padding=2), nn.BatchNorm2d(cout), nn.ReLU(),
                         nn.MaxPool2d(kernel_size=2, stride=2))


class RefConvNet(nn.Module):
    def __init__(self, amp_dtype=None, device_type="cuda"):
        super().__init__()
        self.layer1 = _block(1, 16)
        self.layer2 = _block(16, 32)
        self.fc = nn.Linear(7 * 7 * 32, 10)
        self.amp_dtype, self.device_type = amp_dtype, device_type

    def forward(self, x):
        ctx = (torch.autocast(self.device_type, dtype=self.amp_dtype) if self.amp_dtype is not None
               else contextlib.nullcontext())
        with ctx:
            h = self.layer2(self.layer1(x))
            return self.fc(h.reshape(h.size(0), -1))


def flat_params(model):
    sd = {k.removeprefix("module."): v for k, v in model.state_dict().items()}
    return torch.cat([sd[n].detach().reshape(-1).float() for n, _ in PARAM_SPECS]).cpu()


def flat_grads(model):
    named = {k.removeprefix("module."): v for k, v in model.named_parameters()}
    return torch.cat([named[n].grad.detach().reshape(-1).float() for n, _ in PARAM_SPECS]).cpu()


def bn_buffers(model):
    sd = {k.removeprefix("module."): v for k, v in model.state_dict().items()}
    return torch.cat([sd["layer1.1.running_mean"], sd["layer1.1.running_var"], sd["layer2.1.running_mean"],
                      sd["layer2.1.running_var"]]).float().cpu()


def nbt(model):
    sd = {k.removeprefix("module."): v for k, v in model.state_dict().items()}
    return [int(sd["layer1.1.num_batches_tracked"]), int(sd["layer2.1.num_batches_tracked"])]


class RefTrainer:
    """Loop body of ddp_main.py:86-93 (amp) / origin_main.py:60-66 (fp32) with taps."""

    def __init__(self, model, amp: bool, lr=1e-4, device="cuda"):
        self.model, self.amp = model, amp
        self.criterion = nn.CrossEntropyLoss()
        self.optimizer = torch.optim.SGD(model.parameters(), lr)
        self.scaler = torch.amp.GradScaler(device) if amp else None

    def step(self, images, labels):
        self.model.train()
        outputs = self.model(images)
        loss = self.criterion(outputs, labels)
        self.optimizer.zero_grad()
        rec = {"loss": float(loss.item())}
        if self.amp:
            self.scaler.scale(loss).backward()
            rec["scaled_grads"] = flat_grads(self.model)        # tap T-b
            self.scaler.unscale_(self.optimizer)                # first thing scaler.step does
            g = flat_grads(self.model)                          # tap T-c
            rec["found_inf"] = 0.0 if bool(torch.isfinite(g).all()) else 1.0
            rec["grads"] = g
            self.scaler.step(self.optimizer)
            self.scaler.update()
            rec["scale"] = float(self.scaler.get_scale())
            rec["tracker"] = int(self.scaler._growth_tracker.item())
        else:
            loss.backward()
            rec["grads"] = flat_grads(self.model)
            rec["found_inf"] = 0.0
            self.optimizer.step()
        return rec
