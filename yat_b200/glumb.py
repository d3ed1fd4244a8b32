"""Fused middle of SANA's GLUMBConv (host side of ``yat_glumb_mid_fwd`` / ``yat_glumb_mid_bwd``).

``glumb_mid(u, w_dw, b_dw, H, W)`` maps the token-major output of ``conv_inverted``
(``[B, H*W, 2*Hd]`` bf16) to the input of ``conv_point`` (``[B, H*W, Hd]`` bf16):
SiLU -> depthwise 3x3 (+bias) -> GLU, one kernel forward, one backward; ``u`` and the conv output
are kept for the backward.  Replaces the NCHW cuDNN path of the reference's GLUMBConv
(utils/patch_sana_attention_layers.py:106-111).  No CPU fallback.
"""
import torch

from . import _lib


class _GlumbMid(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, w_dw, b_dw, H, W):
        if not u.is_cuda:
            raise ValueError("yat_b200 ops run on CUDA tensors only (no CPU fallback)")
        if u.dtype != torch.bfloat16 or u.dim() != 3 or u.shape[1] != H * W or u.shape[2] % 2:
            raise ValueError(f"u must be [B, {H * W}, 2*Hd] bf16, got {tuple(u.shape)} {u.dtype}")
        B, _, C2 = u.shape
        Hd = C2 // 2
        if w_dw.numel() != C2 * 9 or b_dw.numel() != C2:
            raise ValueError("depthwise weight must hold [2*Hd, 1, 3, 3] and its bias [2*Hd]")
        u = u.contiguous()
        w = w_dw.detach().to(torch.bfloat16).contiguous()
        bias = b_dw.detach().to(torch.bfloat16).contiguous()
        mid = torch.empty((B, H * W, Hd), dtype=torch.bfloat16, device=u.device)
        need_grad = any(ctx.needs_input_grad[:3])
        c = torch.empty_like(u) if need_grad else None       # depthwise-conv output, kept for the backward
        with _lib.timed("glumb_mid_fwd"):
            rc = _lib.lib().yat_glumb_mid_fwd(_lib.ptr(u), _lib.ptr(w), _lib.ptr(bias), B, H, W, Hd, _lib.ptr(mid),
                                              _lib.ptr(c), _lib.current_stream())
        _lib.check(rc, "yat_glumb_mid_fwd")
        ctx.save_for_backward(u, c, w)
        ctx.hw = (H, W)
        ctx.param_meta = (w_dw.shape, w_dw.dtype, b_dw.shape, b_dw.dtype)
        return mid

    @staticmethod
    def backward(ctx, dmid):
        u, c, w = ctx.saved_tensors
        H, W = ctx.hw
        B, _, C2 = u.shape
        Hd = C2 // 2
        dmid = dmid.contiguous()
        du = torch.empty_like(u)
        dw = torch.empty((C2, 9), dtype=torch.bfloat16, device=u.device)
        db = torch.empty((C2,), dtype=torch.bfloat16, device=u.device)
        ws = torch.empty(_lib.lib().yat_glumb_workspace_bytes(B, H, Hd), dtype=torch.uint8, device=u.device)
        with _lib.timed("glumb_mid_bwd"):
            rc = _lib.lib().yat_glumb_mid_bwd(_lib.ptr(u), _lib.ptr(c), _lib.ptr(w), _lib.ptr(dmid), B, H, W, Hd,
                                              _lib.ptr(du), _lib.ptr(dw), _lib.ptr(db), _lib.ptr(ws),
                                              _lib.current_stream())
        _lib.check(rc, "yat_glumb_mid_bwd")
        w_shape, w_dtype, b_shape, b_dtype = ctx.param_meta
        return du, dw.view(w_shape).to(w_dtype), db.view(b_shape).to(b_dtype), None, None


def glumb_mid(u, w_dw, b_dw, H, W):
    """silu -> depthwise 3x3 (+bias) -> x * silu(gate);  u [B, H*W, 2*Hd] bf16 -> [B, H*W, Hd] bf16."""
    return _GlumbMid.apply(u, w_dw, b_dw, int(H), int(W))
