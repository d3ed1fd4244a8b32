"""ctypes binding of ``libyat_b200.so`` (the C-ABI in ``include/yat_b200.h``).

The library is the only compute path: if it is missing, or a call fails, this module raises —
there is no eager/PyTorch fallback behind any op.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libyat_b200.so")

c_void_p, c_int, c_int64, c_uint64, c_uint32, c_float = (
    ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_float)

# name -> (restype, argtypes): must list every symbol include/yat_b200.h declares
SIGNATURES = {
    "yat_version": (c_int, []),
    "yat_last_error": (ctypes.c_char_p, []),
    "yat_noise_flowmatch": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint64, c_uint64,
                                    c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_void_p, c_void_p]),
    "yat_noise_ddpm": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_uint64, c_uint64, c_int, c_int64,
                               c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_loss_workspace_bytes": (c_int64, []),
    "yat_loss_fused": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_int, c_int,
                               c_uint64, c_uint64, c_int, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p,
                               c_void_p]),
    "yat_repa_workspace_bytes": (c_int64, [c_int, c_int, c_int, c_int]),
    "yat_repa_cos": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_float,
                             c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_linattn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_linattn_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                c_void_p, c_void_p, c_void_p]),
    "yat_linattn_fwd_simt": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_linattn_bwd_simt": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                     c_void_p, c_void_p, c_void_p]),
    "yat_pad_embeddings": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "yat_bucket_assign": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "yat_philox_raw": (c_int, [c_uint64, c_uint64, c_uint32, c_uint32, c_int, c_void_p, c_void_p]),
}

_lib = None


class YatError(RuntimeError):
    pass


def lib():
    """Load (once) and return the ctypes handle; raises if the extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise YatError(
                f"{LIB_PATH} is missing: build it with `python -m yat_b200.build` "
                "(yat_b200 has no CPU or eager fallback)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if a declared symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(status, what):
    """Map a yat_status to a Python exception (ValueError for argument/shape errors)."""
    if status == 0:
        return
    msg = lib().yat_last_error().decode("utf-8", "replace")
    if status in (-1, -2):
        raise ValueError(f"{what}: {msg}")
    raise YatError(f"{what}: {msg} (status {status})")


def ptr(t):
    """Device/host pointer of a tensor, or None."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def current_stream():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
