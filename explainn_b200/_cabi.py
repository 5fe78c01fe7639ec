"""ctypes binding of libexplainn_b200.so (see include/explainn_b200.h).

There is no fallback: if the shared object is missing this module raises at import, and every op
raises if the CUDA call fails.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libexplainn_b200.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        "explainn_b200: %s not found. Build it with `python build_lib.py` "
        "(nvcc, sm_100a); there is no CPU or PyTorch fallback." % LIB_PATH)

lib = C.CDLL(LIB_PATH)

DTYPE_F32, DTYPE_F16, DTYPE_F32_PREACT = 0, 1, 2
LOSS_BCE, LOSS_MSE = 0, 1

_P, _LL, _I, _F, _ULL = C.c_void_p, C.c_longlong, C.c_int, C.c_float, C.c_ulonglong

# name -> argtypes (restype int unless listed in _RESTYPES)
SIGNATURES = {
    "explainn_last_error": [],
    "explainn_abi_version": [],
    "explainn_num_sms": [],
    "explainn_pack_onehot": [_P, _I, _LL, _I, _LL, _P, _P, _P, _P],
    "explainn_conv_tables_floats": [_I, _I],
    "explainn_conv_wq_floats": [_I, _I],
    "explainn_conv_prepare": [_P, _P, _P, _P, _P, _P, _F, _I, _I, _P, _P, _P, _P, _P],
    "explainn_conv_pool_fwd": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P],
    "explainn_conv_pool_fwd_train": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "explainn_conv_tc_image_halfs": [_I, _I],
    "explainn_conv_tc_prepare": [_P, _P, _P, _P, _P, _P, _F, _P, _I, _I, _P, _P, _P],
    "explainn_conv_pool_fwd_tc": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P],
    "explainn_conv_pool_fwd_tc_fp16": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P],
    "explainn_conv_pool_fwd_train_tc": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P],
    "explainn_unit_activations": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _I, _P],
    "explainn_interp_unit_max_scratch_floats": [_I],
    "explainn_interp_unit_max": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P, _I, _P],
    "explainn_interp_unit_max_noalloc": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _I, _P],
    "explainn_interp_thresholds": [_P, _I, _P, _P, _P, _P, _P],
    "explainn_interp_pwm_count_scratch_bytes": [_LL, _I, _I, _I],
    "explainn_interp_pwm_count": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P],
    "explainn_interp_tc_bounds": [_P, _P, _P, _I, _I, _P, _P],
    "explainn_interp_tc_list_bytes": [_I],
    "explainn_interp_unit_max_tc": [_P, _P, _LL, _I, _I, _I] + [_P] * 7 + [_LL, _P, _P],
    "explainn_interp_pwm_count_tc": [_P, _P, _LL, _I, _I, _I] + [_P] * 13 + [_LL, _P, _P],
    "explainn_interp_seq_argmax": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "explainn_act_unit_max": [_P, _I, _LL, _I, _I, _P, _P],
    "explainn_act_thresholds": [_P, _I, _I, _P, _P, _P],
    "explainn_act_pwm_count": [_P, _I, _P, _P, _LL, _I, _I, _I, _P, _P, _P, _P, _P],
    "explainn_revcomp_packed": [_P, _P, _LL, _I, _P, _P, _P],
    "explainn_encode_dna": [_P, _P, _LL, _I, _P, _P, _P],
    "explainn_unpack_onehot": [_P, _P, _LL, _I, _P, _I, _P],
    "explainn_pwm_score": [_P, _P, _LL, _I, _I, _I, _P, _P, _P, _I, _I, _P, _P],
    "explainn_fc_stack_fwd_eval": [_P, _LL, _I, _I] + [_P] * 12 + [_F, _P, _P],
    "explainn_linear_fwd": [_P, _LL, _I, _I, _P, _P, _P, _P],
    "explainn_fc1_stack_fwd_eval": [_P, _LL, _I, _I] + [_P] * 6 + [_F, _P, _P],
    "explainn_fc1_stack_fwd_train": [_P, _LL, _I, _I, _I] + [_P] * 6 + [_F, _F] + [_P] * 8 + [_P],
    "explainn_fc1_stack_bwd": [_P, _LL, _I, _I, _I] + [_P] * 8 + [_P] * 9 + [_P],
    "explainn_input_cooc_stats": [_P, _P, _LL, _I, _I, _P, _P],
    "explainn_bn1_batch_stats": [_P, _P, _P, _I, _I, _LL, _F, _P, _P, _P, _P, _P, _P, _P],
    "explainn_fc_nchunk": [_LL],
    "explainn_fc_stack_fwd_train": [_P, _LL, _I, _I, _I] + [_P] * 12 + [_F, _F, _F, _ULL, _P, _P] + [_P] * 12 + [_P],
    "explainn_fc_stack_bwd": [_P, _LL, _I, _I, _I] + [_P] * 14 + [_F, _ULL, _P, _P] + [_P] * 14 + [_P],
    "explainn_fc_stack_fwd_train_phase": [_I, _LL, _P, _LL, _I, _I, _I] + [_P] * 12 + [_F, _F, _F, _ULL, _P, _P] + [_P] * 12 + [_P],
    "explainn_fc_stack_bwd_phase": [_I, _P, _LL, _I, _I, _I] + [_P] * 14 + [_F, _ULL, _P, _P] + [_P] * 14 + [_P],
    "explainn_conv_pool_bwd_phase": [_I, _LL, _P, _P, _LL, _I, _I, _I] + [_P] * 10 + [_F, _P, _I] + [_P] * 4 + [_P],
    "explainn_conv_bwd_nsplit": [_I],
    "explainn_conv_pool_bwd": [_P, _P, _LL, _I, _I, _I] + [_P] * 10 + [_F, _P, _I] + [_P] * 4 + [_P],
    "explainn_pool_general": [_P, _LL, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P],
    "explainn_conv_pool_bwd_general": [_P, _P, _LL, _I, _I, _I, _I, _I] + [_P] * 10 + [_F, _P, _I] + [_P] * 4 + [_P],
    "explainn_linear_bwd": [_P, _P, _P, _LL, _I, _I, _P, _P, _P, _P],
    "explainn_hidden_fwd_train": [_P, _LL, _I, _I] + [_P] * 6 + [_F, _F, _F, _ULL, _P] + [_P] * 4 + [_P],
    "explainn_hidden_fwd_eval": [_P, _LL, _I, _I] + [_P] * 6 + [_F] + [_P] * 3 + [_P],
    "explainn_hidden_bwd": [_P, _LL, _I, _I] + [_P] * 7 + [_F, _ULL, _P] + [_P] * 7 + [_P],
    "explainn_loss_scratch_bytes": [],
    "explainn_loss_fwd_bwd": [_P, _P, _LL, _I, _I, _F, _P, _P, _P, _P],
    "explainn_adam_step": [_P, _P, _P, _P, _LL, _P, _F, _F, _F, _F, _F, _P],
    "explainn_reduce_partials": [_P, _P, _LL, _I, _P],
    "explainn_counters_inc": [_P, _P, _P, _P, _P],
    "explainn_p2p_flag_bytes": [],
    "explainn_p2p_alloc": [_LL, _P],
    "explainn_p2p_free": [_P],
    "explainn_p2p_export": [_P, _P],
    "explainn_p2p_open": [_P, _P],
    "explainn_p2p_close": [_P],
    "explainn_p2p_error": [_P, _P],
    "explainn_p2p_adam_step": [_P, _P, _P, _LL, _LL, _P, _I, _P, _P, _I, _I, _ULL, _F, _F, _F, _F, C.c_double, _P],
}
_RESTYPES = {
    "explainn_last_error": C.c_char_p,
    "explainn_conv_tables_floats": _LL,
    "explainn_conv_wq_floats": _LL,
    "explainn_interp_unit_max_scratch_floats": _LL,
    "explainn_interp_pwm_count_scratch_bytes": _LL,
    "explainn_interp_tc_list_bytes": _LL,
    "explainn_loss_scratch_bytes": _LL,
    "explainn_conv_tc_image_halfs": _LL,
    "explainn_p2p_flag_bytes": _LL,
}


def _bind():
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the library lacks a declared symbol
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, C.c_int)


def register(name, argtypes, restype=C.c_int):
    """Bind one more symbol (used by modules that add entry points)."""
    SIGNATURES[name] = argtypes
    if restype is not C.c_int:
        _RESTYPES[name] = restype
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype


_bind()


def last_error():
    return lib.explainn_last_error().decode()


def ptr(t):
    """Device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise ValueError("explainn_b200: expected a CUDA tensor, got device %s (no CPU path exists)" % t.device)
    if not t.is_contiguous():
        raise ValueError("explainn_b200: tensor must be contiguous")
    return t.data_ptr()


def stream():
    return torch.cuda.current_stream().cuda_stream


def on_device_of(argnum=1):
    """Decorator: run the wrapped method with the device of its tensor argument current, so that the stream handed
    to the kernels (torch's current stream of the CURRENT device) and per-device caches belong to the GPU that holds
    the data -- a model on cuda:1 works while cuda:0 is current, as with the reference's plain `device` argument."""
    import functools

    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*args, **kw):
            t = args[argnum] if len(args) > argnum else None
            if isinstance(t, torch.Tensor) and t.is_cuda and t.device.index != torch.cuda.current_device():
                with torch.cuda.device(t.device):
                    return fn(*args, **kw)
            return fn(*args, **kw)
        return wrapper
    return deco


LAUNCHES = 0   # kernels launched through this binding (bench.py reports it)
_LAUNCH_COST = {"explainn_interp_unit_max": 3, "explainn_interp_unit_max_noalloc": 3, "explainn_interp_pwm_count": 3, "explainn_interp_unit_max_tc": 4, "explainn_interp_pwm_count_tc": 4, "explainn_input_cooc_stats": 1, "explainn_fc_stack_fwd_train": 3,
                "explainn_fc_stack_bwd": 6, "explainn_conv_pool_bwd": 2, "explainn_adam_step": 2}


def call(name, *args):
    global LAUNCHES
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = "%s failed (code %d): %s" % (name, rc, last_error())
        if rc in (1, 3, 4):
            raise ValueError(msg)
        raise RuntimeError(msg)
    LAUNCHES += _LAUNCH_COST.get(name, 1)
    return rc
