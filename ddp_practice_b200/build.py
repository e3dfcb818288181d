"""Build libddpstep.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB_DIR = os.path.join(PKG, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libddpstep.so")
SOURCES = ["api.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]


def _newest_source_mtime():
    m = 0.0
    for root in (CSRC, os.path.join(PKG, "..", "include")):
        for f in os.listdir(root):
            if f.endswith((".cu", ".cuh", ".h")):
                m = max(m, os.path.getmtime(os.path.join(root, f)))
    return m


LAST_ACTION = "not built in this process"


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA library if it is missing or older than its sources; LAST_ACTION says which happened."""
    global LAST_ACTION
    if not force and os.path.exists(LIB_PATH) and os.path.getmtime(LIB_PATH) >= _newest_source_mtime():
        LAST_ACTION = "reused (library newer than every source by mtime)"
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, *NVCC_FLAGS, "-o", LIB_PATH, *[os.path.join(CSRC, s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed ({' '.join(cmd)}):\n{res.stdout}\n{res.stderr}")
    if verbose:
        print(res.stderr)
    LAST_ACTION = "compiled with nvcc for sm_100a"
    return LIB_PATH


def build_tools() -> None:
    """Microbenchmarks and the tracing build of the library under tools/ (not tracked; they travel to the GPU box)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    tools = os.path.join(PKG, "..", "tools")
    arch = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17"]
    for src in ("microbench", "microbench_step"):
        subprocess.run([nvcc, *arch, "-o", os.path.join(tools, src + ".bin"), os.path.join(tools, src + ".cu")], check=True)
    subprocess.run([nvcc, *arch, "-Xcompiler", "-fPIC", "-shared", "-DDDPSTEP_TRACE", "-o",
                    os.path.join(tools, "libddpstep_trace.so"), os.path.join(CSRC, "api.cu")], check=True)


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
