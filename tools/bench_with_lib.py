"""Run bench.py against another build of the library (A/B of two library versions on the same box):
python tools/bench_with_lib.py tools/libddpstep_prev.so --steps 20 --warmup 5 ..."""
import os, runpy, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ddp_practice_b200 import build, _lib
so = os.path.abspath(sys.argv[1])
build.LIB_PATH = so; _lib.LIB_PATH = so
import ctypes
have = ctypes.CDLL(so)
_lib.SYMBOLS = [s for s in _lib.SYMBOLS if hasattr(have, s[0])]
sys.argv = [os.path.join(ROOT, "bench.py")] + sys.argv[2:]
runpy.run_path(sys.argv[0], run_name="__main__")
