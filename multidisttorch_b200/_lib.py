"""ctypes binding of include/mdt_b200.h (the same stub INTEGRATION.md shows).
There is no fallback: if the shared library is missing, loading fails loudly."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libmdt_b200.so")

MDT_GEMM = {"fp32": 0, "tf32": 1, "tf32x3": 2, "bf16": 3}
MDT_ACT = {"none": 0, "relu": 1, "sigmoid": 2}
MDT_EXCHANGE = {"auto": 0, "nccl": 1, "p2p": 2, "p2p_two_shot": 3}
SEGMENTS = ("W1", "b1", "W2x", "b2x", "W3", "b3", "W4", "b4")


class mdt_cfg(C.Structure):
    _fields_ = [("in_dim", C.c_int32), ("hidden", C.c_int32), ("latent", C.c_int32), ("max_batch", C.c_int32),
                ("gemm_mode", C.c_int32), ("use_graph", C.c_int32), ("exchange", C.c_int32), ("reserved", C.c_int32)]


class mdt_adam(C.Structure):
    _fields_ = [("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


class mdt_sgd(C.Structure):
    _fields_ = [("lr", C.c_double), ("momentum", C.c_double), ("dampening", C.c_double), ("weight_decay", C.c_double),
                ("nesterov", C.c_int32), ("reserved", C.c_int32)]


class mdt_batch_source(C.Structure):
    _fields_ = [("x_pool", C.c_void_p), ("images_u8", C.c_void_p), ("indices", C.c_void_p), ("n_batches", C.c_int64),
                ("n_images", C.c_int64)]


_P = C.c_void_p
_PROTOS = {
    "mdt_version": (C.c_int, []),
    "mdt_last_error": (C.c_char_p, []),
    "mdt_param_layout": (C.c_int, [C.POINTER(mdt_cfg), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "mdt_engine_create": (C.c_int, [C.POINTER(mdt_cfg), _P, C.c_int, C.c_int, _P, C.POINTER(_P)]),
    "mdt_engine_destroy": (None, [_P]),
    "mdt_engine_stream": (C.c_int, [_P, C.POINTER(_P)]),
    "mdt_engine_bind_buffers": (C.c_int, [_P, _P, _P, _P, _P, C.c_size_t]),
    "mdt_engine_broadcast_params": (C.c_int, [_P]),
    "mdt_engine_set_adam": (C.c_int, [_P, C.POINTER(mdt_adam), C.c_int64]),
    "mdt_engine_set_rng": (C.c_int, [_P, C.c_uint64, C.c_uint64]),
    "mdt_engine_set_sgd": (C.c_int, [_P, C.POINTER(mdt_sgd), C.c_int64]),
    "mdt_train_steps": (C.c_int, [_P, C.POINTER(mdt_batch_source), C.c_int, C.c_int64, C.c_int, _P]),
    "mdt_engine_exchange_bench": (C.c_int, [_P, C.c_int, C.POINTER(C.c_double)]),
    "mdt_engine_get_state": (C.c_int, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]),
    "mdt_engine_input_buffer": (C.c_int, [_P, C.POINTER(_P)]),
    "mdt_forward": (C.c_int, [_P, _P, C.c_int, _P, _P, _P, _P]),
    "mdt_backward": (C.c_int, [_P, _P, _P, _P]),
    "mdt_allreduce_grads": (C.c_int, [_P]),
    "mdt_adam_step": (C.c_int, [_P, C.c_float]),
    "mdt_train_step": (C.c_int, [_P, _P, C.c_int, _P, _P]),
    "mdt_train_step_many": (C.c_int, [C.POINTER(_P), C.POINTER(_P), C.c_int, C.c_int]),
    "mdt_group_create": (C.c_int, [C.POINTER(_P), C.c_int, C.POINTER(_P)]),
    "mdt_group_train_steps": (C.c_int, [_P, C.POINTER(mdt_batch_source), C.c_int, C.c_int64, C.c_int]),
    "mdt_group_destroy": (None, [_P]),
    "mdt_train_step_host": (C.c_int, [_P, _P, C.c_int, _P, C.POINTER(C.c_float)]),
    "mdt_engine_set_sm_share": (C.c_int, [_P, C.c_int]),
    "mdt_train_step_host_submit": (C.c_int, [_P, _P, C.c_int, _P, C.POINTER(C.c_uint32)]),
    "mdt_train_step_host_wait": (C.c_int, [_P, C.c_uint32, C.POINTER(C.c_float)]),
    "mdt_train_steps_host_submit": (C.c_int, [_P, _P, C.c_int, C.c_int, C.POINTER(C.c_uint32)]),
    "mdt_eval_step": (C.c_int, [_P, _P, C.c_int, _P, _P, _P]),
    "mdt_decode": (C.c_int, [_P, _P, C.c_int, _P]),
    "mdt_engine_workspace": (C.c_int, [_P, C.c_char_p, C.POINTER(_P), C.POINTER(C.c_int)]),
    "mdt_engine_launch_count": (C.c_int, [_P, C.POINTER(C.c_int64)]),
    "mdt_engine_exchange_handle": (C.c_int, [_P, _P]),
    "mdt_engine_set_peers": (C.c_int, [_P, _P, C.c_int]),
    "mdt_engine_disable_peers": (C.c_int, [_P]),
    "mdt_engine_exchange_status": (C.c_int, [_P, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "mdt_engine_profile_step": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_char_p, C.c_size_t]),
    "mdt_adam_flat": (C.c_int, [_P, _P, _P, _P, C.c_size_t, C.POINTER(mdt_adam), C.c_int64, C.c_float, _P]),
    "mdt_reparam": (C.c_int, [_P, _P, C.c_int, _P, C.c_uint64, C.c_uint64, C.c_int, C.c_int, _P, C.c_int, _P, _P]),
    "mdt_loss_bce_kl": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, _P]),
    "mdt_loss_bce_kl_noalloc": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, C.c_int, _P, _P, _P, _P, _P]),
    "mdt_gather_u8": (C.c_int, [_P, C.c_int64, _P, C.c_int, C.c_int, _P, _P]),
    "mdt_linear": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _P, _P]),
    "mdt_allreduce_bench": (C.c_int, [_P, _P, C.c_size_t, C.c_int, _P, C.POINTER(C.c_double)]),
}

_lib = None


def exported_symbols():
    return tuple(_PROTOS)


def load():
    """dlopen the C-ABI library and attach prototypes.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA extension is not built (run `make -C multidisttorch_b200/csrc` "
            "or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)          # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class MdtError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise MdtError(f"mdt error {rc}: {load().mdt_last_error().decode(errors='replace')}")
