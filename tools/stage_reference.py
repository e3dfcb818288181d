#!/usr/bin/env python
"""Stage the reference's own scripts for a run on the GPU box (where /root/reference does not exist).

Run HERE (build container) before `gpurun`:   python tools/stage_reference.py

Writes only into the git-ignored scratch directory baseline/_ref/ (it travels with the gpurun snapshot, it never
enters history):

  ref_ddp_spawn.py / ref_ddp_torchrun.py     byte-for-byte copies of /root/reference/ddp_main.py / ddp_main_torchrun.py
  shim_ddp_spawn.py / shim_ddp_torchrun.py   the same files after EXACTLY the two import edits of INTEGRATION.md section 1
                                             (applied with str.replace, asserted to hit once each)
  data/MNIST/raw/*-ubyte                     synthetic idx files (torchvision only checks that they exist,
                                             torchvision/datasets/mnist.py:168-172): 12,800 train / 2,000 test images,
                                             class-dependent templates + noise so that an epoch visibly learns
  seed/sitecustomize.py                      `torch.manual_seed(3407)` at interpreter start-up: the scripts leave the
                                             model init unseeded (DDP broadcasts rank 0), and two runs can only be
                                             compared from the same init.  The scripts themselves are not touched.

tests/test_gpu_reference_scripts.py runs the pairs when this directory exists."""
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
OUT = os.path.join(ROOT, "baseline", "_ref")

EDITS = [("import torch.nn as nn\n", "from ddp_practice_b200.shim import nn\n"),
         ("from torch.cuda.amp import GradScaler\n", "from ddp_practice_b200.shim import GradScaler\n")]


def idx_images(path, a):
    with open(path, "wb") as f:
        f.write(struct.pack(">IIII", 0x00000803, a.shape[0], 28, 28))
        f.write(a.astype(np.uint8).tobytes())


def idx_labels(path, a):
    with open(path, "wb") as f:
        f.write(struct.pack(">II", 0x00000801, a.shape[0]))
        f.write(a.astype(np.uint8).tobytes())


def synth(n, rng, templates):
    y = rng.integers(0, 10, n)
    noise = rng.integers(0, 256, (n, 28, 28)).astype(np.float32)
    x = np.clip(0.55 * templates[y] + 0.45 * noise, 0, 255).round()
    return x, y


def main():
    if not os.path.isdir(REF):
        sys.exit(f"{REF} is not mounted here: staging happens in the build container")
    os.makedirs(OUT, exist_ok=True)
    for src, name in (("ddp_main.py", "ddp_spawn"), ("ddp_main_torchrun.py", "ddp_torchrun")):
        text = open(os.path.join(REF, src)).read()
        open(os.path.join(OUT, f"ref_{name}.py"), "w").write(text)
        edited = text
        for old, new in EDITS:
            assert edited.count(old) == 1, (src, old)
            edited = edited.replace(old, new)
        open(os.path.join(OUT, f"shim_{name}.py"), "w").write(edited)
    raw = os.path.join(OUT, "data", "MNIST", "raw")
    os.makedirs(raw, exist_ok=True)
    rng = np.random.default_rng(3407)
    templates = rng.integers(0, 256, (10, 28, 28)).astype(np.float32)
    xtr, ytr = synth(12800, rng, templates)
    xte, yte = synth(2000, rng, templates)
    idx_images(os.path.join(raw, "train-images-idx3-ubyte"), xtr)
    idx_labels(os.path.join(raw, "train-labels-idx1-ubyte"), ytr)
    idx_images(os.path.join(raw, "t10k-images-idx3-ubyte"), xte)
    idx_labels(os.path.join(raw, "t10k-labels-idx1-ubyte"), yte)
    os.makedirs(os.path.join(OUT, "seed"), exist_ok=True)
    open(os.path.join(OUT, "seed", "sitecustomize.py"), "w").write(
        "# every process of a staged run starts from the same RNG state (see tools/stage_reference.py)\n"
        "import torch\ntorch.manual_seed(3407)\n")
    print(f"staged into {OUT}: {sorted(os.listdir(OUT))}")


if __name__ == "__main__":
    main()
