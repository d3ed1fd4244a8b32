"""Fused middle of SANA's GLUMBConv (host side of ``yat_glumb_mid_fwd`` / ``yat_glumb_mid_bwd``).

``glumb_mid(u, w_dw, b_dw, H, W)`` maps the token-major output of ``conv_inverted``
(``[B, H*W, 2*Hd]`` bf16) to the input of ``conv_point`` (``[B, H*W, Hd]`` bf16):
SiLU -> depthwise 3x3 (+bias) -> GLU, one kernel forward, one backward; ``u`` and the conv output
are kept for the backward.  Replaces the NCHW cuDNN path of the reference's GLUMBConv
(utils/patch_sana_attention_layers.py:106-111).  No CPU fallback.
"""
import torch
import torch.nn.functional as F

from . import _lib


def _fwd(u, w_dw, b_dw, H, W, need_grad):
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
    c = torch.empty_like(u) if need_grad else None           # depthwise-conv output, kept for the backward
    with _lib.timed("glumb_mid_fwd"):
        rc = _lib.lib().yat_glumb_mid_fwd(_lib.ptr(u), _lib.ptr(w), _lib.ptr(bias), B, H, W, Hd, _lib.ptr(mid),
                                          _lib.ptr(c), _lib.current_stream())
    _lib.check(rc, "yat_glumb_mid_fwd")
    return u, c, w, mid


def _bwd(u, c, w, dmid, H, W, want_colsum):
    """-> du, dw [2*Hd, 9], db [2*Hd], colsum(du) [2*Hd] or None (all bf16)."""
    B, _, C2 = u.shape
    Hd = C2 // 2
    dmid = dmid.contiguous()
    du = torch.empty_like(u)
    dw = torch.empty((C2, 9), dtype=torch.bfloat16, device=u.device)
    db = torch.empty((C2,), dtype=torch.bfloat16, device=u.device)
    dbu = torch.empty((C2,), dtype=torch.bfloat16, device=u.device) if want_colsum else None
    ws = torch.empty(_lib.lib().yat_glumb_workspace_bytes(B, H, Hd), dtype=torch.uint8, device=u.device)
    with _lib.timed("glumb_mid_bwd"):
        rc = _lib.lib().yat_glumb_mid_bwd(_lib.ptr(u), _lib.ptr(c), _lib.ptr(w), _lib.ptr(dmid), B, H, W, Hd,
                                          _lib.ptr(du), _lib.ptr(dw), _lib.ptr(db), _lib.ptr(dbu), _lib.ptr(ws),
                                          _lib.current_stream())
    _lib.check(rc, "yat_glumb_mid_bwd")
    return du, dw, db, dbu


class _GlumbMid(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u, w_dw, b_dw, H, W):
        u, c, w, mid = _fwd(u, w_dw, b_dw, H, W, any(ctx.needs_input_grad[:3]))
        ctx.save_for_backward(u, c, w)
        ctx.hw = (H, W)
        ctx.param_meta = (w_dw.shape, w_dw.dtype, b_dw.shape, b_dw.dtype)
        return mid

    @staticmethod
    def backward(ctx, dmid):
        u, c, w = ctx.saved_tensors
        H, W = ctx.hw
        du, dw, db, _ = _bwd(u, c, w, dmid, H, W, False)
        w_shape, w_dtype, b_shape, b_dtype = ctx.param_meta
        return du, dw.view(w_shape).to(w_dtype), db.view(b_shape).to(b_dtype), None, None


class _InvertedGlumbMid(torch.autograd.Function):
    """conv_inverted (1x1 = GEMM with the bias in its epilogue) + the fused middle as one autograd node: the bias
    gradient of conv_inverted is the column sum of du, which the backward kernel forms from its fp32 values while
    writing du - autograd's separate ``du.sum(0)`` pass over [B*H*W, 2*Hd] (367 MB at SANA-1.6B) disappears."""

    @staticmethod
    def forward(ctx, x, w_inv, b_inv, w_dw, b_dw, H, W):
        u = F.linear(x, w_inv, b_inv)
        u, c, w, mid = _fwd(u, w_dw, b_dw, H, W, any(ctx.needs_input_grad[:5]))
        ctx.save_for_backward(x, w_inv, u, c, w)
        ctx.hw = (H, W)
        ctx.param_meta = (w_dw.shape, w_dw.dtype, b_dw.shape, b_dw.dtype, b_inv.dtype)
        return mid

    @staticmethod
    def backward(ctx, dmid):
        x, w_inv, u, c, w = ctx.saved_tensors
        H, W = ctx.hw
        need = ctx.needs_input_grad
        du, dw, db, dbu = _bwd(u, c, w, dmid, H, W, need[2])
        w_shape, w_dtype, b_shape, b_dtype, bi_dtype = ctx.param_meta
        du2 = du.view(-1, du.shape[-1])
        dx = (du2 @ w_inv).view_as(x) if need[0] else None
        dw_inv = du2.t() @ x.reshape(-1, x.shape[-1]) if need[1] else None
        return (dx, dw_inv, dbu.to(bi_dtype) if need[2] else None, dw.view(w_shape).to(w_dtype) if need[3] else None,
                db.view(b_shape).to(b_dtype) if need[4] else None, None, None)


def glumb_mid(u, w_dw, b_dw, H, W):
    """silu -> depthwise 3x3 (+bias) -> x * silu(gate);  u [B, H*W, 2*Hd] bf16 -> [B, H*W, Hd] bf16."""
    return _GlumbMid.apply(u, w_dw, b_dw, int(H), int(W))


def inverted_glumb_mid(x, w_inv, b_inv, w_dw, b_dw, H, W):
    """``glumb_mid(F.linear(x, w_inv, b_inv), ...)`` with conv_inverted's bias gradient taken from the backward
    kernel; x [B, H*W, C] bf16, w_inv [2*Hd, C], b_inv [2*Hd]."""
    return _InvertedGlumbMid.apply(x, w_inv, b_inv, w_dw, b_dw, int(H), int(W))
