"""Host-side logic of the drop-in shim that needs no GPU."""
import ast
import os

import pytest
import torch
import torch.nn as tnn

from ddp_practice_b200 import shim
from tests.refmodel import RefConvNet

REF = "/root/reference"
EDITS = {
    "import torch.nn as nn": "from ddp_practice_b200.shim import nn",
    "from torch.cuda.amp import GradScaler": "from ddp_practice_b200.shim import GradScaler",
}


def test_nn_namespace_forwards_to_torch_nn_except_the_three_overrides():
    assert shim.nn.Conv2d is tnn.Conv2d and shim.nn.Module is tnn.Module and shim.nn.Sequential is tnn.Sequential
    assert shim.nn.SyncBatchNorm is shim.SyncBatchNorm
    assert shim.nn.CrossEntropyLoss is shim.CrossEntropyLoss
    assert shim.nn.parallel.DistributedDataParallel is shim.DistributedDataParallel


def test_only_the_reference_convnet_is_accepted():
    shim._check_convnet(RefConvNet(None, "cpu"))
    shim._check_convnet(shim.SyncBatchNorm.convert_sync_batchnorm(RefConvNet(torch.float16, "cpu")))
    bad = RefConvNet(None, "cpu")
    bad.fc = tnn.Linear(1568, 11)
    with pytest.raises(NotImplementedError):
        shim._check_convnet(bad)
    with pytest.raises(NotImplementedError):
        shim._check_convnet(tnn.Linear(4, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        shim.DistributedDataParallel(RefConvNet(None, "cpu"))  # no CPU path


def test_grad_scaler_defaults_and_state_dict_keys():
    s = shim.GradScaler()
    sd = s.state_dict()  # grad_scaler.py:645-651
    assert sd == {"scale": 65536.0, "growth_factor": 2.0, "backoff_factor": 0.5, "growth_interval": 2000,
                  "_growth_tracker": 0}
    s.load_state_dict({**sd, "scale": 1024.0, "_growth_tracker": 7})
    assert s.get_scale() == 1024.0 and s.state_dict()["_growth_tracker"] == 7


def test_cross_entropy_passthrough_on_real_tensors():
    c = shim.CrossEntropyLoss()
    lg, y = torch.randn(4, 10), torch.tensor([1, 2, 3, 4])
    assert torch.equal(c(lg, y), tnn.CrossEntropyLoss()(lg, y))


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference not mounted (GPU box)")
@pytest.mark.parametrize("script", ["ddp_main.py", "ddp_main_torchrun.py"])
def test_reference_scripts_need_only_two_import_edits(script):
    """north_star: the demo scripts run with import-level changes only.  Apply the two edits to
    the reference source, check nothing else had to change, and that every name the script
    takes from `nn` / GradScaler resolves in the shim."""
    src = open(os.path.join(REF, script)).read()
    out = src
    for old, new in EDITS.items():
        assert out.count(old) == 1
        out = out.replace(old, new)
    changed = [(a, b) for a, b in zip(src.splitlines(), out.splitlines()) if a != b]
    assert len(changed) == 2 and len(src.splitlines()) == len(out.splitlines())
    tree = ast.parse(out)
    used = {n.attr for n in ast.walk(tree) if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name) and n.value.id == "nn"}
    for name in used:
        assert getattr(shim.nn, name) is not None
    assert {"SyncBatchNorm", "CrossEntropyLoss", "parallel", "Conv2d", "BatchNorm2d"} <= used
    methods = {n.attr for n in ast.walk(tree) if isinstance(n, ast.Attribute) and isinstance(n.value, ast.Name) and n.value.id == "scaler"}
    assert methods == {"scale", "step", "update", "state_dict"}
    for m in methods:
        assert callable(getattr(shim.GradScaler, m))


def test_sgd_hyper_parameters_are_read_from_the_param_group():
    """the shim hands torch.optim.SGD's knobs to the fused tail kernel; unsupported ones fail loudly
    before anything is launched."""
    p = [tnn.Parameter(torch.zeros(3))]
    g = torch.optim.SGD(p, 1e-2, momentum=0.9, weight_decay=5e-4, nesterov=True).param_groups[0]
    assert shim._sgd_knobs(g) == dict(momentum=0.9, weight_decay=5e-4, dampening=0.0, nesterov=True)
    assert shim._sgd_knobs(torch.optim.SGD(p, 1e-4).param_groups[0]) == dict(momentum=0.0, weight_decay=0.0,
                                                                            dampening=0.0, nesterov=False)
    with pytest.raises(NotImplementedError, match="maximize"):
        shim.GradScaler().step(torch.optim.SGD(p, 1e-4, maximize=True))
    # several param groups are supported (ddpstep_set_sgd_group); an optimizer over parameters that belong to no shim
    # model is refused before anything is launched
    two = torch.optim.SGD([{"params": p}, {"params": [tnn.Parameter(torch.zeros(2))]}], 1e-4)
    with pytest.raises(RuntimeError, match="does not hold the parameters"):
        shim.GradScaler().step(two)


def test_config_struct_mirrors_the_header():
    """ctypes Config / Ptrs must list exactly the fields of ddpstep_config / ddpstep_ptrs in include/ddpstep.h, in order."""
    import re
    from ddp_practice_b200 import _lib
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "ddpstep.h")).read()

    def fields(struct):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct, struct), src, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            rest = decl.split(None, 1)[1]  # drop the type
            names += [n.strip().lstrip("*") for n in rest.split(",")]
        return names

    assert fields("ddpstep_config") == [n for n, _ in _lib.Config._fields_]
    assert fields("ddpstep_ptrs") == [n for n, _ in _lib.Ptrs._fields_]
    cfg = _lib.Config()
    _lib.load().ddpstep_default_config(cfg)
    assert cfg.abi_version == 3 and cfg.momentum == 0.0 and cfg.nesterov == 0 and abs(cfg.lr - 1e-4) < 1e-10


def test_c_abi_argument_errors_need_no_gpu():
    """Every entry point validates its arguments before it touches CUDA, so the error paths can be exercised here:
    bad world size / ABI version / optimizer knobs at create time, and the allocation-free entry point's null,
    alignment and size checks (its message names the 1024-image limit)."""
    import ctypes as C
    from ddp_practice_b200 import _lib
    lib = _lib.load()
    cfg = _lib.Config()
    lib.ddpstep_default_config(C.byref(cfg))
    ctx = C.c_void_p()
    cfg.world, cfg.rank = 9, 0
    assert lib.ddpstep_create(C.byref(cfg), C.byref(ctx)) == -1 and b"world" in lib.ddpstep_last_error(None)
    lib.ddpstep_default_config(C.byref(cfg))
    cfg.abi_version = 1
    assert lib.ddpstep_create(C.byref(cfg), C.byref(ctx)) == -1 and b"ABI" in lib.ddpstep_last_error(None)
    lib.ddpstep_default_config(C.byref(cfg))
    cfg.nesterov = 1  # without momentum: torch/optim/sgd.py:68 raises too
    assert lib.ddpstep_create(C.byref(cfg), C.byref(ctx)) == -1 and b"nesterov" in lib.ddpstep_last_error(None)
    lib.ddpstep_default_config(C.byref(cfg))
    cfg.batch_cap = 0
    assert lib.ddpstep_create(C.byref(cfg), C.byref(ctx)) == -1
    # allocation-free entry point: fake but well-formed addresses, nothing is dereferenced before the checks
    a16 = 0x10000
    call = lib.ddpstep_conv1_bnstats_noalloc
    assert call(None, a16, 8, 1, 1, 1e-5, a16, a16, None) == -1 and b"null" in lib.ddpstep_last_error(None)
    assert call(a16, a16, 8, 7, 1, 1e-5, a16, a16, None) == -1 and b"dtype" in lib.ddpstep_last_error(None)
    assert call(a16, a16 + 4, 8, 1, 1, 1e-5, a16, a16, None) == -1 and b"aligned" in lib.ddpstep_last_error(None)
    assert call(a16, a16, 8, 1, 1, 1e-5, a16, a16 + 2, None) == -1 and b"aligned" in lib.ddpstep_last_error(None)
    assert call(a16, a16, 1025, 1, 1, 1e-5, a16, a16, None) == -1 and b"1024" in lib.ddpstep_last_error(None)
    assert call(a16, a16, 0, 1, 1, 1e-5, a16, a16, None) == -1
    # null contexts are refused, not dereferenced
    assert lib.ddpstep_train_step(None, a16, a16, 8, None) == -1
    assert lib.ddpstep_set_sgd_group(None, 1, 1e-3, 0.0, 0.0, 0.0, 0, None) == -1
    assert lib.ddpstep_allreduce_bucket(None, a16, 8, 1, None) == -1
    assert lib.ddpstep_launch_count(None) == 0 and lib.ddpstep_destroy(None) == 0
    assert lib.ddpstep_kernel_name(14) == b"step" and lib.ddpstep_kernel_name(99) == b""
