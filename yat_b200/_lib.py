"""ctypes binding of ``libyat_b200.so`` (the C-ABI in ``include/yat_b200.h``).

The library is the only compute path: if it is missing, or a call fails, this module raises —
there is no eager/PyTorch fallback behind any op.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libyat_b200.so")

c_void_p, c_int, c_int64, c_uint64, c_uint32, c_float, c_double = (
    ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_float, ctypes.c_double)

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
    "yat_block_workspace_bytes": (c_int64, [c_int, c_int, c_int]),
    "yat_ln_modulate_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_float, c_void_p,
                                    c_void_p, c_void_p, c_void_p]),
    "yat_ln_modulate_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int, c_int,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_add_ln_modulate_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_float,
                                        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_ln_modulate_bwd_res": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int,
                                        c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_gate_residual_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_gate_residual_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p,
                                      c_void_p, c_void_p]),
    "yat_glumb_workspace_bytes": (c_int64, [c_int, c_int, c_int]),
    "yat_glumb_mid_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                  c_void_p]),
    "yat_glumb_mid_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_colsum_workspace_bytes": (c_int64, [c_int64, c_int]),
    "yat_colsum_bf16": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p]),
    "yat_colsum_static_max_bytes": (c_int64, []),
    "yat_colsum_bf16_static": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p]),
    "yat_optim_chunk_elems": (c_int, []),
    "yat_optim_workspace_bytes": (c_int64, [c_int]),
    "yat_grad_clip_coef": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p]),
    "yat_adamw_step": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_double, c_double, c_double,
                               c_double, c_double, c_int64, c_void_p]),
    "yat_adamw_ema_step": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_double, c_double, c_double,
                                   c_double, c_double, c_int64, c_void_p, c_double, c_void_p]),
    "yat_ema_step": (c_int, [c_void_p, c_void_p, c_int, c_int, c_double, c_void_p]),
    "yat_pad_embeddings": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "yat_linattn_fwd_ld": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_linattn_bwd_ld": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                   c_void_p, c_void_p, c_int64, c_void_p]),
    "yat_head_pad": (c_int, [c_void_p, c_void_p, c_float, c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_head_unpad": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_gemm_bf16_tn": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int64, c_int64, c_int64,
                                 c_int, c_void_p]),
    "yat_qkv_linattn_state_bytes": (c_int64, [c_int, c_int, c_int]),
    "yat_qkv_linattn_fwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "yat_linattn_fwd_state": (c_int, [c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "yat_xattn_max_keys": (c_int, []),
    "yat_xattn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int64, c_int64,
                              c_float, c_void_p, c_void_p, c_void_p]),
    "yat_xattn_bwd_workspace_bytes": (c_int64, [c_int, c_int, c_int]),
    "yat_xattn_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                              c_int, c_int64, c_int64, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "yat_bucket_assign": (c_int, [c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "yat_philox_raw": (c_int, [c_uint64, c_uint64, c_uint32, c_uint32, c_int, c_void_p, c_void_p]),
}

_lib = None

# C-ABI calls that launch at least one of our kernels, counted per entry point (bench.py reports the
# sum over its timed region as `gpu_launches`).
LAUNCHING = ("yat_noise_flowmatch", "yat_noise_ddpm", "yat_loss_fused", "yat_repa_cos", "yat_linattn_fwd",
             "yat_linattn_bwd", "yat_linattn_fwd_simt", "yat_linattn_bwd_simt", "yat_pad_embeddings",
             "yat_philox_raw", "yat_ln_modulate_fwd", "yat_ln_modulate_bwd", "yat_add_ln_modulate_fwd",
             "yat_ln_modulate_bwd_res", "yat_gate_residual_fwd",
             "yat_gate_residual_bwd", "yat_glumb_mid_fwd", "yat_glumb_mid_bwd", "yat_grad_clip_coef",
             "yat_adamw_step", "yat_adamw_ema_step", "yat_ema_step", "yat_colsum_bf16", "yat_colsum_bf16_static", "yat_head_pad", "yat_head_unpad",
             "yat_linattn_fwd_ld", "yat_linattn_bwd_ld", "yat_gemm_bf16_tn", "yat_qkv_linattn_fwd",
             "yat_linattn_fwd_state", "yat_xattn_fwd", "yat_xattn_bwd")
launch_counts = {name: 0 for name in LAUNCHING}


def total_launches():
    return sum(launch_counts.values())


class _Counted:
    """ctypes function + a call counter (a Python int increment; no device work)."""
    __slots__ = ("fn", "name")

    def __init__(self, fn, name):
        self.fn, self.name = fn, name

    def __call__(self, *args):
        launch_counts[self.name] += 1
        return self.fn(*args)


class _Handle:
    pass


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
        cdll = ctypes.CDLL(LIB_PATH)
        handle = _Handle()
        handle.cdll = cdll
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(cdll, name)  # AttributeError if a declared symbol is not exported
            fn.restype = res
            fn.argtypes = args
            setattr(handle, name, _Counted(fn, name) if name in LAUNCHING else fn)
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


# Optional in-situ kernel timing: bench.py sets ``timing_sink`` to a list and every op wrapper then
# brackets its C-ABI call with CUDA events recorded on the launching (current) stream.
timing_sink = None


class timed:
    __slots__ = ("name", "e0")

    def __init__(self, name):
        self.name = name
        self.e0 = None

    def __enter__(self):
        if timing_sink is not None:
            import torch
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if self.e0 is not None:
            import torch
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            timing_sink.append((self.name, self.e0, e1))
        return False
