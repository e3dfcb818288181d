"""B200-native fused DDP + AMP training step of the gbbin/DDP-practice ConvNet.

`Engine` is the thin host wrapper over the C ABI (include/ddpstep.h); `shim` mirrors the
PyTorch call surface `ddp_main.py` uses so the demo runs with import-level changes only."""
from .engine import Engine, NativeError  # noqa: F401
