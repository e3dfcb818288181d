"""ctypes binding of include/ddpstep.h.  There is no CPU or PyTorch fallback: if the CUDA
library is missing or does not load, importing the engine fails loudly."""
import ctypes as C
import os

from .build import LIB_PATH

N_PARAMS = 29034
N_BNBUF = 96
HANDLE_BYTES = 64
FP32, FP16, BF16 = 0, 1, 2
DTYPES = {"fp32": FP32, "fp16": FP16, "bf16": BF16}


class Config(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("device", C.c_int32),
        ("dtype", C.c_int32), ("batch_cap", C.c_int32), ("use_scaler", C.c_int32), ("sync_bn", C.c_int32),
        ("use_graph", C.c_int32), ("conv_bias_double_round", C.c_int32),
        ("lr", C.c_float), ("init_scale", C.c_float), ("growth_factor", C.c_float), ("backoff_factor", C.c_float),
        ("growth_interval", C.c_int32), ("bn_momentum", C.c_float), ("bn_eps", C.c_float),
        ("momentum", C.c_float), ("weight_decay", C.c_float), ("dampening", C.c_float), ("nesterov", C.c_int32),
    ]


class Ptrs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "params", "grads", "grads_scaled", "bn_buffers", "num_batches_tracked", "scale", "growth_tracker",
        "found_inf", "loss", "logits", "counts", "input_stage", "label_stage", "momentum_buffer", "momentum_ready")]


# every symbol include/ddpstep.h declares: (name, restype, argtypes)
_P = C.c_void_p
SYMBOLS = [
    ("ddpstep_default_config", None, [C.POINTER(Config)]),
    ("ddpstep_create", C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    ("ddpstep_destroy", C.c_int, [_P]),
    ("ddpstep_last_error", C.c_char_p, [_P]),
    ("ddpstep_export_handle", C.c_int, [_P, _P]),
    ("ddpstep_import_peers", C.c_int, [_P, _P, C.c_int32]),
    ("ddpstep_load_state", C.c_int, [_P, _P, _P, _P]),
    ("ddpstep_read_state", C.c_int, [_P, _P, _P, _P, _P, _P]),
    ("ddpstep_set_scaler", C.c_int, [_P, C.c_float, C.c_int32]),
    ("ddpstep_state_ptrs", C.c_int, [_P, C.POINTER(Ptrs)]),
    ("ddpstep_train_step", C.c_int, [_P, _P, _P, C.c_int32, _P]),
    ("ddpstep_train_steps", C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, _P]),
    ("ddpstep_train_steps_indexed", C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, _P]),
    ("ddpstep_train_step_accumulate", C.c_int, [_P, _P, _P, C.c_int32, _P]),
    ("ddpstep_train_step_host", C.c_int, [_P, _P, _P, C.c_int32, _P, _P]),
    ("ddpstep_train_step_indexed", C.c_int, [_P, _P, _P, _P, C.c_int32, _P]),
    ("ddpstep_eval_batch_indexed", C.c_int, [_P, _P, _P, _P, C.c_int32, _P]),
    ("ddpstep_eval_batch", C.c_int, [_P, _P, _P, C.c_int32, _P]),
    ("ddpstep_reset_counts", C.c_int, [_P, _P]),
    ("ddpstep_reduce_counts", C.c_int, [_P, _P]),
    ("ddpstep_read_counts", C.c_int, [_P, _P, _P, _P]),
    ("ddpstep_read_activation", C.c_int, [_P, C.c_int32, _P, C.c_int64]),
    ("ddpstep_profile_step", C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, _P, C.c_int32, _P]),
    ("ddpstep_kernel_name", C.c_char_p, [C.c_int32]),
    ("ddpstep_conv1_bnstats_noalloc", C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_float, _P, _P, _P]),
    ("ddpstep_set_sgd", C.c_int, [_P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int32, _P]),
    ("ddpstep_set_sgd_group", C.c_int, [_P, C.c_uint32, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int32, _P]),
    ("ddpstep_set_peer_timeout_ms", C.c_int, [_P, C.c_double]),
    ("ddpstep_allreduce_only", C.c_int, [_P, _P]),
    ("ddpstep_allreduce_bucket", C.c_int, [_P, _P, C.c_int64, C.c_int32, _P]),
    ("ddpstep_launch_count", C.c_int64, [_P]),
    ("ddpstep_device_status", C.c_int, [_P]),
]

_lib = None


def load():
    """dlopen the in-tree library and type its entry points."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it first (python -c 'import __graft_entry__ as g; g.build()'). "
            "ddp_practice_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, res, args in SYMBOLS:
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
