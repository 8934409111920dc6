"""ctypes binding of libb200sde.so (the C ABI declared in include/b200_sde.h).

There is NO CPU fallback: if the library cannot be loaded, or a tensor is not on a CUDA device, the ops raise.
"""
import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200sde.so")
_lock = threading.Lock()
_lib = None

c_int, c_ll, c_float, c_u64, c_vp = ctypes.c_int, ctypes.c_longlong, ctypes.c_float, ctypes.c_uint64, ctypes.c_void_p

# name -> argtypes (restype is int unless listed in _RESTYPES)
_SIGNATURES = {
    "b200_abi_version": [],
    "b200_build_info": [],
    "b200_upfirdn2d_out_size": [c_int] * 6,
    "b200_upfirdn2d_f32": [c_vp, c_vp, c_vp] + [c_int] * 13 + [c_vp],
    "b200_upfirdn2d_nhwc": [c_vp, c_vp, c_vp] + [c_int] * 11 + [c_vp],
    "b200_fused_bias_act_f32": [c_vp, c_vp, c_vp, c_vp, c_ll, c_ll, c_int, c_int, c_int, c_float, c_float, c_vp],
    "b200_ns_cip_advect": [c_vp, c_vp, c_vp, c_ll, c_ll, c_int, c_int, c_int, c_float, c_float, c_vp],
    "b200_ns_update_density": [c_vp, c_vp, c_vp, c_int, c_int, c_int, c_float, c_float, c_vp],
    "b200_ns_velocity_nonadvect": [c_vp, c_vp, c_vp, c_int, c_int, c_int, c_float, c_float, c_vp],
    "b200_ns_update_velocity": [c_vp, c_vp, c_vp, c_vp, c_int, c_int, c_int, c_float, c_float, c_int, c_vp],
    "b200_ns_update_pressure": [c_vp, c_vp, c_vp, c_int, c_int, c_int, c_float, c_float, c_vp],
    "b200_ns_step_fused": [c_vp] * 6 + [c_int, c_int, c_int, c_float, c_float, c_vp],
    "b200_rk_stage": [c_vp, c_vp, c_vp, c_vp, c_ll, c_int, c_vp, ctypes.c_double, c_ll, c_vp],
    "b200_rk_error": [c_vp, c_vp, c_vp, c_vp, c_ll, c_int, c_vp, ctypes.c_double, ctypes.c_double, ctypes.c_double, c_ll,
                      c_vp],
    "b200_sample_sumsq": [c_vp, c_vp, c_vp, c_vp, c_int, c_ll, c_u64, c_u64, c_ll, c_vp, c_vp],
    "b200_langevin_norm_sums": [c_vp, c_vp, c_vp, c_int, c_vp],
    "b200_langevin_norms_static_max_batch": [],
    "b200_xchg_max_ranks": [],
    "b200_xchg_alloc": [ctypes.POINTER(c_vp), ctypes.c_char_p],
    "b200_xchg_open": [ctypes.c_char_p, ctypes.POINTER(c_vp)],
    "b200_xchg_close": [c_vp],
    "b200_xchg_free": [c_vp],
    "b200_langevin_norm_sums_xchg": [c_vp, c_vp, c_vp, c_int, ctypes.POINTER(c_vp), c_int, c_int, c_vp],
    "b200_langevin_norms_static": [c_vp, c_vp, c_vp, c_int, c_ll, c_u64, c_u64, c_ll, c_vp, c_vp],
    "b200_sde_update": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_int, c_vp, c_float, c_float, c_int, c_ll,
                        c_u64, c_u64, c_ll, c_vp],
    "b200_fill_normal": [c_vp, c_ll, c_float, c_u64, c_u64, c_ll, c_vp],
    "b200_mask_mix": [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, ctypes.POINTER(c_float), ctypes.POINTER(c_float),
                      c_int, c_int, c_ll, c_u64, c_u64, c_ll, c_vp],
    "b200_step_advance": [c_vp, c_int, c_vp],
    "b200_grid_sample_fwd": [c_int, c_vp, c_vp, c_vp] + [c_int] * 10 + [c_vp],
    "b200_grid_sample_bwd": [c_int, c_vp, c_vp, c_vp, c_vp, c_vp] + [c_int] * 10 + [c_vp],
    "b200_grid_sample_grad2": [c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp] + [c_int] * 10 + [c_vp],
}
_RESTYPES = {"b200_build_info": ctypes.c_char_p}


def register_signature(name, argtypes, restype=None):
    """Lets the other host modules declare the entry points they bind (engine, grid_sample, ...)."""
    _SIGNATURES[name] = argtypes
    if restype is not None:
        _RESTYPES[name] = restype
    if _lib is not None:
        fn = getattr(_lib, name)
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, c_int)


def lib():
    """Loads the library once.  Raises RuntimeError (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libb200sde.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` or "
                "`python b_pinn_kalman_filter_b200/build.py`. There is no CPU fallback." % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        for name, argtypes in _SIGNATURES.items():
            fn = getattr(l, name)  # AttributeError if the .so is stale: loud on purpose
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, c_int)
        _lib = l
    return _lib


_ERR = {-1: "bad argument", -2: "unsupported shape", -3: "CUDA driver entry point unavailable",
        -4: "size beyond the static scratch capacity of an allocation-free entry point"}


def check(rc, what):
    if rc == 0:
        return
    if rc <= -1000:
        raise RuntimeError("%s: CUDA launch failed (cudaError %d)" % (what, -rc - 1000))
    raise RuntimeError("%s: %s (code %d)" % (what, _ERR.get(rc, "error"), rc))


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and t.device.type != "cuda":
            raise RuntimeError("b200 ops need CUDA tensors (got device '%s'); there is no CPU fallback" % t.device)


def ptr(t):
    return None if t is None else c_vp(t.data_ptr())


def stream():
    return c_vp(torch.cuda.current_stream().cuda_stream)
