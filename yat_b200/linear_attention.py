"""SANA ReLU linear attention op (host side of ``yat_linattn_fwd`` / ``yat_linattn_bwd``).

``sana_linear_attention(q, k, v, heads)`` takes the outputs of ``to_q/to_k/to_v``
([B, N, heads*32] bf16) and returns the attention output in the same layout — the part of
diffusers' ``SanaLinearAttnProcessor2_0`` between the projections and ``to_out``.  Forward and
backward are tcgen05/TMEM kernels; the backward recomputes K^T V (nothing but q, k, v is saved,
which is also what gradient checkpointing — train_sana.py:63 — wants).
"""
import torch

from . import _lib


def _launch_fwd(q, k, v, heads, impl):
    B, N, C = q.shape
    o = torch.empty_like(q)
    fn = _lib.lib().yat_linattn_fwd if impl == "sm100" else _lib.lib().yat_linattn_fwd_simt
    with _lib.timed("linattn_fwd_" + impl):
        rc = fn(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), B, N, heads, C // heads, _lib.ptr(o), _lib.current_stream())
    _lib.check(rc, "yat_linattn_fwd")
    return o


def _launch_bwd(q, k, v, do, heads, impl):
    B, N, C = q.shape
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    fn = _lib.lib().yat_linattn_bwd if impl == "sm100" else _lib.lib().yat_linattn_bwd_simt
    with _lib.timed("linattn_bwd_" + impl):
        rc = fn(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(do), B, N, heads, C // heads,
                _lib.ptr(dq), _lib.ptr(dk), _lib.ptr(dv), _lib.current_stream())
    _lib.check(rc, "yat_linattn_bwd")
    return dq, dk, dv


class _SanaLinearAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, heads, impl):
        for t in (q, k, v):
            if not t.is_cuda:
                raise ValueError("yat_b200 ops run on CUDA tensors only (no CPU fallback)")
            if t.dtype != torch.bfloat16:
                raise ValueError("linear attention expects bf16 q, k, v")
        if q.shape != k.shape or q.shape != v.shape or q.dim() != 3 or q.shape[-1] % heads:
            raise ValueError(f"bad q/k/v shapes {tuple(q.shape)} {tuple(k.shape)} {tuple(v.shape)} for {heads} heads")
        q, k, v = q.contiguous(), k.contiguous(), v.contiguous()
        ctx.save_for_backward(q, k, v)
        ctx.heads, ctx.impl = heads, impl
        return _launch_fwd(q, k, v, heads, impl)

    @staticmethod
    def backward(ctx, do):
        q, k, v = ctx.saved_tensors
        dq, dk, dv = _launch_bwd(q, k, v, do.contiguous(), ctx.heads, ctx.impl)
        return dq, dk, dv, None, None


def sana_linear_attention(q, k, v, heads, impl="sm100"):
    """impl: "sm100" = tensor-core kernels (the product path); "simt" = the CUDA-core fp32
    restatement kept for GPU-side cross-checks in tests."""
    if impl not in ("sm100", "simt"):
        raise ValueError(f"unknown impl {impl!r}")
    return _SanaLinearAttention.apply(q, k, v, heads, impl)


class _SanaLinearAttentionQKV(torch.autograd.Function):
    """Same kernels on the three column thirds of ONE fused projection output qkv [B, N, 3*C] (row pitch 3*C): no
    q / k / v copies, and the backward writes dq | dk | dv straight into the gradient of that projection, so its
    input gradient is a single GEMM over K = 3*C instead of three GEMMs and two accumulation passes."""

    @staticmethod
    def forward(ctx, qkv, heads):
        if not qkv.is_cuda or qkv.dtype != torch.bfloat16 or qkv.dim() != 3 or qkv.shape[-1] % (3 * heads):
            raise ValueError("fused linear attention expects a CUDA bf16 [B, N, 3*heads*D] tensor (no CPU fallback)")
        qkv = qkv.contiguous()
        B, N, C3 = qkv.shape
        C = C3 // 3
        o = torch.empty((B, N, C), dtype=qkv.dtype, device=qkv.device)
        base = qkv.data_ptr()
        with _lib.timed("linattn_fwd_sm100"):
            rc = _lib.lib().yat_linattn_fwd_ld(base, base + 2 * C, base + 4 * C, C3, B, N, heads, C // heads,
                                               _lib.ptr(o), _lib.current_stream())
        _lib.check(rc, "yat_linattn_fwd_ld")
        ctx.save_for_backward(qkv)
        ctx.heads = heads
        return o

    @staticmethod
    def backward(ctx, do):
        (qkv,) = ctx.saved_tensors
        B, N, C3 = qkv.shape
        C = C3 // 3
        do = do.contiguous()
        dqkv = torch.empty_like(qkv)
        base, gbase = qkv.data_ptr(), dqkv.data_ptr()
        with _lib.timed("linattn_bwd_sm100"):
            rc = _lib.lib().yat_linattn_bwd_ld(base, base + 2 * C, base + 4 * C, C3, _lib.ptr(do), B, N, ctx.heads,
                                               C // ctx.heads, gbase, gbase + 2 * C, gbase + 4 * C, C3,
                                               _lib.current_stream())
        _lib.check(rc, "yat_linattn_bwd_ld")
        return dqkv, None


def sana_linear_attention_qkv(qkv, heads):
    """qkv [B, N, 3*heads*32] bf16 = [q | k | v] along the last axis (one fused projection) -> [B, N, heads*32]."""
    return _SanaLinearAttentionQKV.apply(qkv, int(heads))


class _SanaLinearAttentionProj(torch.autograd.Function):
    """x -> to_q | to_k | to_v -> ReLU linear attention, with the projection on the repo's own tcgen05 GEMM and the
    attention's phase 1 (relu(k)^T v, sum relu(k)) as that GEMM's epilogue (``yat_qkv_linattn_fwd``); the attention
    forward then streams only q and the 33 x 32 per-head state (``yat_linattn_fwd_state``).  SURVEY.md §8(f) N1.
    Backward: the same tcgen05 backward kernel on the saved qkv, then the two projection-gradient GEMMs."""

    @staticmethod
    def forward(ctx, x, w_qkv, heads):
        if not (x.is_cuda and w_qkv.is_cuda) or x.dtype != torch.bfloat16 or w_qkv.dtype != torch.bfloat16:
            raise ValueError("fused projection + linear attention expects CUDA bf16 tensors (no CPU fallback)")
        if x.dim() != 3 or w_qkv.dim() != 2 or w_qkv.shape[1] != x.shape[-1] or w_qkv.shape[0] % (3 * heads):
            raise ValueError(f"bad shapes x {tuple(x.shape)} w_qkv {tuple(w_qkv.shape)} for {heads} heads")
        x, w_qkv = x.contiguous(), w_qkv.contiguous()
        B, N, Cin = x.shape
        C = w_qkv.shape[0] // 3
        L = _lib.lib()
        qkv = torch.empty((B, N, 3 * C), dtype=torch.bfloat16, device=x.device)
        state = torch.empty(L.yat_qkv_linattn_state_bytes(B, N, heads) // 4, dtype=torch.float32, device=x.device)
        o = torch.empty((B, N, C), dtype=torch.bfloat16, device=x.device)
        with _lib.timed("qkv_linattn_fwd"):
            rc = L.yat_qkv_linattn_fwd(_lib.ptr(x), _lib.ptr(w_qkv), B, N, heads, C // heads, Cin, _lib.ptr(qkv),
                                       _lib.ptr(state), _lib.current_stream())
        _lib.check(rc, "yat_qkv_linattn_fwd")
        with _lib.timed("linattn_fwd_state"):
            rc = L.yat_linattn_fwd_state(_lib.ptr(qkv), 3 * C, _lib.ptr(state), B, N, heads, C // heads, _lib.ptr(o),
                                         _lib.current_stream())
        _lib.check(rc, "yat_linattn_fwd_state")
        ctx.save_for_backward(x, w_qkv, qkv)
        ctx.heads = heads
        return o

    @staticmethod
    def backward(ctx, do):
        x, w_qkv, qkv = ctx.saved_tensors
        B, N, C3 = qkv.shape
        C = C3 // 3
        do = do.contiguous()
        dqkv = torch.empty_like(qkv)
        base, gbase = qkv.data_ptr(), dqkv.data_ptr()
        with _lib.timed("linattn_bwd_sm100"):
            rc = _lib.lib().yat_linattn_bwd_ld(base, base + 2 * C, base + 4 * C, C3, _lib.ptr(do), B, N, ctx.heads,
                                               C // ctx.heads, gbase, gbase + 2 * C, gbase + 4 * C, C3,
                                               _lib.current_stream())
        _lib.check(rc, "yat_linattn_bwd_ld")
        g2 = dqkv.view(B * N, C3)
        dx = (g2 @ w_qkv).view_as(x) if ctx.needs_input_grad[0] else None
        dw = g2.t() @ x.view(B * N, -1) if ctx.needs_input_grad[1] else None
        return dx, dw, None


def sana_linear_attention_proj(x, w_qkv, heads):
    """x [B, N, Cin] bf16, w_qkv [3*heads*32, Cin] bf16 (to_q | to_k | to_v weights stacked) -> [B, N, heads*32]."""
    return _SanaLinearAttentionProj.apply(x, w_qkv, int(heads))
