"""Build the library with -DDDPSTEP_TRACE into a scratch .so and print in-kernel phase timestamps (CTA 0)."""
import ctypes as C, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
so = os.path.join(ROOT, "tools", "libddpstep_trace.so")
from ddp_practice_b200 import build, _lib
build.LIB_PATH = so; _lib.LIB_PATH = so
from ddp_practice_b200 import Engine
from oracle import convnet_oracle as O
from tests.refmodel import RefConvNet
lib = _lib.load()
lib.ddpstep_read_trace.argtypes = [C.c_void_p]
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
torch.manual_seed(0)
st = O.OracleState.from_state_dict(RefConvNet(None, "cpu").state_dict())
e = Engine(dtype="fp16", batch_cap=B)
bn = torch.cat([st.running_mean[0], st.running_var[0], st.running_mean[1], st.running_var[1]])
e.load_state(st.flat_params().numpy(), bn.numpy())
x, y = O.synthetic_mnist(B); x, y = x.cuda(), y.cuda()
for i in range(20): e.train_step(x, y)
buf = (C.c_ulonglong * 64)()
lib.ddpstep_read_trace(buf)
t = list(buf)
names = sys.argv[2].split(",") if len(sys.argv) > 2 else [str(i) for i in range(16)]
base = t[0]
for i, n in enumerate(names):
    print(f"{n:28s} +{(t[i]-base)/1e3:8.2f} us" + (f"  (d {(t[i]-t[i-1])/1e3:6.2f})" if i else ""))
