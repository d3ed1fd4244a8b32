"""Attention module and processors (diffusers attention-processor protocol).

The reference installs processors with ``attn.set_processor(...)`` and calls them as
``processor(attn, hidden_states, encoder_hidden_states=None, attention_mask=None)``
(utils/patch_sana_attention_layers.py:125-137).  ``Attention`` below carries the attribute names
processors and PEFT configs rely on (``to_q to_k to_v to_out.0``, ``heads``, ``norm_q``,
``norm_k``); ``SanaLinearAttnProcessorB200`` is the slot the tcgen05 kernel drops into, and
``AttnProcessor2_0`` (softmax SDPA, library call) is what ``patch_sana_attention_layers`` swaps in
for ``modified_blocks`` and what cross-attention uses.
"""
import os
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from . import _lib
from .linear import Linear
from .linear_attention import sana_linear_attention, sana_linear_attention_proj, sana_linear_attention_qkv


class SanaLinearAttnProcessorB200:
    """Drop-in for diffusers ``SanaLinearAttnProcessor2_0``: projections stay ``attn.to_q/k/v``
    (possibly PEFT-wrapped), the ReLU linear-attention core runs as one tcgen05 kernel."""

    def __init__(self, impl: str = "sm100", fuse_qkv: bool = True, fuse_projection: bool = False):
        self.impl = impl
        self.fuse_qkv = fuse_qkv
        # fuse_projection: run to_q | to_k | to_v on the repo's own tcgen05 GEMM with the attention's phase 1 as its
        # epilogue (yat_qkv_linattn_fwd).  Off by default: measured on B200 at the SANA-1.6B layer
        # size the pair (fused GEMM 400 us + state attention 55 us) is 17 % slower than (library GEMM 337 us +
        # attention 61 us) - the operand staging of the epilogue costs the GEMM its sixth pipeline stage.
        self.fuse_projection = fuse_projection

    def _can_fuse_qkv(self, attn, hidden_states, encoder_hidden_states):
        if not self.fuse_qkv or self.impl != "sm100" or encoder_hidden_states is not hidden_states:
            return False
        mods = (attn.to_q, attn.to_k, attn.to_v)
        if any(hasattr(m, "adapter") or m._forward_hooks or m._forward_pre_hooks for m in mods):
            return False                                   # LoRA / LoKr wraps and REPA-style hooks see the modules
        if len({m.bias is None for m in mods}) != 1 or len({m.weight.shape for m in mods}) != 1:
            return False
        return hidden_states.is_cuda and hidden_states.dtype == torch.bfloat16 and attn.to_q.weight.dtype == torch.bfloat16

    def __call__(self, attn, hidden_states, encoder_hidden_states=None, attention_mask=None):
        original_dtype = hidden_states.dtype
        if encoder_hidden_states is None:
            encoder_hidden_states = hidden_states
        if attn.norm_q is not None or attn.norm_k is not None:
            raise ValueError("qk_norm is not supported by the B200 linear-attention kernel "
                             "(SANA-1.6B uses qk_norm=None, utils/patched_sana_transformer.py:109)")
        if self._can_fuse_qkv(attn, hidden_states, encoder_hidden_states):
            # ONE projection GEMM on the concatenated weights (parameters stay separate, HF layout): the kernels read
            # q | k | v as column thirds of its output and write dq | dk | dv into its gradient, so the input gradient
            # is a single GEMM over K = 3*C instead of three GEMMs and two accumulation passes
            w = torch.cat([attn.to_q.weight, attn.to_k.weight, attn.to_v.weight], dim=0)
            b = None if attn.to_q.bias is None else torch.cat([attn.to_q.bias, attn.to_k.bias, attn.to_v.bias], dim=0)
            if self.fuse_projection and b is None and hidden_states.dim() == 3:
                out = sana_linear_attention_proj(hidden_states, w, attn.heads)
            else:
                out = sana_linear_attention_qkv(F.linear(hidden_states, w, b), attn.heads)
            out = attn.to_out[0](out.to(original_dtype))
            return attn.to_out[1](out)
        query = attn.to_q(hidden_states)
        key = attn.to_k(encoder_hidden_states)
        value = attn.to_v(encoder_hidden_states)
        out = sana_linear_attention(query.to(torch.bfloat16), key.to(torch.bfloat16), value.to(torch.bfloat16),
                                    attn.heads, impl=self.impl)
        out = out.to(original_dtype)
        out = attn.to_out[0](out)
        out = attn.to_out[1](out)
        return out


class _HeadPad(torch.autograd.Function):
    """x [B, T, H*d] -> [B, T, H, da] with column d = tail (constant or one bf16 value per token), rest 0."""

    @staticmethod
    def forward(ctx, x, H, d, da, tail_const, tail_row):
        B, T, _ = x.shape
        x = x.contiguous()
        out = torch.empty((B, T, H, da), dtype=x.dtype, device=x.device)
        with _lib.timed("head_pad"):
            rc = _lib.lib().yat_head_pad(_lib.ptr(x), _lib.ptr(tail_row), float(tail_const), B * T, H, d, da,
                                         _lib.ptr(out), _lib.current_stream())
        _lib.check(rc, "yat_head_pad")
        ctx.dims = (H, d)
        return out

    @staticmethod
    def backward(ctx, g):
        H, d = ctx.dims
        return _unpad(g, d), None, None, None, None, None


class _HeadUnpad(torch.autograd.Function):
    """xa [B, T, H, da] (any batch / token / head strides, unit stride along the head dim) -> [B, T, H*d]."""

    @staticmethod
    def forward(ctx, xa, d):
        ctx.dims = (xa.shape[2], d, xa.shape[3])
        return _unpad(xa, d)

    @staticmethod
    def backward(ctx, g):
        H, d, da = ctx.dims
        return _HeadPad.apply(g, H, d, da, 0.0, None), None


def _unpad(xa, d):
    B, T, H, _ = xa.shape
    if xa.stride(3) != 1 or any(st % 8 for st in xa.stride()[:3]) or xa.data_ptr() % 16:
        xa = xa.contiguous()
    out = torch.empty((B, T, H * d), dtype=xa.dtype, device=xa.device)
    with _lib.timed("head_unpad"):
        rc = _lib.lib().yat_head_unpad(_lib.ptr(xa), xa.stride(0), xa.stride(1), xa.stride(2), B, T, H, d,
                                       _lib.ptr(out), _lib.current_stream())
    _lib.check(rc, "yat_head_unpad")
    return out


_key_tail_cache = {}


def _key_tail(bias, d):
    """bias [B, 1, 1, L] (additive, -10000 on padded keys) -> bf16 [B*L] = bias * sqrt(d); one entry cached so the
    20 blocks of a forward that share the mask tensor build it once."""
    key = (bias.data_ptr(), bias._version, bias.numel(), d)
    hit = _key_tail_cache.get("k")
    if hit is not None and hit[0] == key:
        return hit[2]
    tail = (bias.reshape(-1).float() * (d ** 0.5)).to(torch.bfloat16)
    _key_tail_cache["k"] = (key, bias, tail)           # holding `bias` keeps its storage (and the pointer) alive
    return tail


def _sdpa_key_bias_folded(q, k, v, bias, H, d):
    """SDPA with a per-key additive bias (the -10000 text-padding mask) folded into one extra head column:
    q' = [q, 1, 0..], k' = [k, bias_j * sqrt(d), 0..], v unchanged  =>  scale * q'.k' = scale * q.k + bias_j,
    so the library kernel runs WITHOUT a mask tensor.  Measured on B200 for the SANA-1.6B cross-attention
    ([16,20,1024,112] x [16,20,320,112], scripts/bench_xattn2.py): an additive mask takes cuDNN's SDPA from 67 to
    207 us forward and from 392 to 757 us forward + backward.  The padded copies are own streaming kernels
    (yat_head_pad / yat_head_unpad, each the other's backward).  Masked keys still weigh exactly 0 after the softmax
    (exp(-10000) underflows in fp32 either way).  q [B, N, H*d], k / v [B, L, H*d] token-major, bias [B, 1, 1, L];
    returns [B, N, H*d]."""
    da = (d + 1 + 15) // 16 * 16
    B = q.shape[0]
    qa = _HeadPad.apply(q, H, d, da, 1.0, None)
    ka = _HeadPad.apply(k, H, d, da, 0.0, _key_tail(bias, d))
    # only q and k carry the extra column: the library takes a value head dim (d) different from the q / k one (da)
    # at the same speed, and returns the output token-major, so v, o and their gradients are never copied
    out = F.scaled_dot_product_attention(qa.transpose(1, 2), ka.transpose(1, 2), v.view(B, -1, H, d).transpose(1, 2),
                                         dropout_p=0.0, is_causal=False, scale=d ** -0.5)
    return out.transpose(1, 2).reshape(B, -1, H * d)


_bias_f32_cache = {}


def _bias_f32(bias, B, L):
    """Additive key mask -> fp32 [B, L]; one entry cached so the 20 blocks of a pass that share the mask convert it once."""
    if bias is None:
        return None
    key = (bias.data_ptr(), bias._version, bias.numel())
    hit = _bias_f32_cache.get("b")
    if hit is not None and hit[0] == key:
        return hit[2]
    bias_f = bias.reshape(B, L).float().contiguous()
    _bias_f32_cache["b"] = (key, bias, bias_f)      # holding `bias` keeps its storage (and the pointer) alive
    return bias_f


def _xattn_own_forward(q, k, v, bias, H, d):
    """Forward-only cross-attention on the own tcgen05 kernel (``yat_xattn_fwd``, csrc/xattn_sm100.cu): no padded copies, the
    key bias read as it is.  q [B, N, H*d], k / v [B, L, H*d] bf16 token-major (row-pitched views allowed), bias [B, L]-shaped
    additive mask or None.  There is no own backward, so the caller uses this only when no gradient is needed."""
    B, N, _ = q.shape
    L = k.shape[1]
    bias_f = _bias_f32(bias, B, L)
    out = torch.empty((B, N, H * d), dtype=torch.bfloat16, device=q.device)
    with _lib.timed("xattn_fwd"):
        rc = _lib.lib().yat_xattn_fwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(bias_f), B, H, N, L, d, q.stride(1),
                                      k.stride(1), float(d) ** -0.5, _lib.ptr(out), None, _lib.current_stream())
    _lib.check(rc, "yat_xattn_fwd")
    return out


class _XAttnOwn(torch.autograd.Function):
    """Cross-attention with both directions on the own tcgen05 kernels (yat_xattn_fwd / yat_xattn_bwd): the training path of
    attn2 (``AttnProcessor2_0.own_backward``, default on; 408 against 478 us per SANA-1.6B layer)."""

    @staticmethod
    def forward(ctx, q, k, v, bias_f, H):
        B, N, C = q.shape
        L, d = k.shape[1], C // H
        out = torch.empty((B, N, C), dtype=torch.bfloat16, device=q.device)
        lse = torch.empty((B, H, N), dtype=torch.float32, device=q.device)
        with _lib.timed("xattn_fwd"):
            rc = _lib.lib().yat_xattn_fwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(bias_f), B, H, N, L, d, q.stride(1),
                                          k.stride(1), float(d) ** -0.5, _lib.ptr(out), _lib.ptr(lse), _lib.current_stream())
        _lib.check(rc, "yat_xattn_fwd")
        ctx.save_for_backward(q, k, v, bias_f, out, lse)
        ctx.H = H
        return out

    @staticmethod
    def backward(ctx, do):
        q, k, v, bias_f, out, lse = ctx.saved_tensors
        H = ctx.H
        B, N, C = q.shape
        L, d = k.shape[1], C // H
        do = do.contiguous()
        dq = torch.empty((B, N, C), dtype=torch.bfloat16, device=q.device)
        dk = torch.empty((B, L, C), dtype=torch.bfloat16, device=q.device)
        dv = torch.empty((B, L, C), dtype=torch.bfloat16, device=q.device)
        lib = _lib.lib()
        ws = torch.empty(lib.yat_xattn_bwd_workspace_bytes(B, H, N), dtype=torch.uint8, device=q.device)
        with _lib.timed("xattn_bwd"):
            rc = lib.yat_xattn_bwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(bias_f), _lib.ptr(out), _lib.ptr(lse),
                                   _lib.ptr(do), B, H, N, L, d, q.stride(1), k.stride(1), float(d) ** -0.5, _lib.ptr(dq),
                                   _lib.ptr(dk), _lib.ptr(dv), _lib.ptr(ws), _lib.current_stream())
        _lib.check(rc, "yat_xattn_bwd")
        return dq, dk, dv, None, None


def _xattn_layout_ok(q, k, v, d):
    return (k.shape[1] <= _lib.lib().yat_xattn_max_keys() and d <= 128
            and q.stride(2) == 1 and k.stride(2) == 1 and v.stride(2) == 1 and k.stride(1) == v.stride(1)
            and q.stride(0) == q.shape[1] * q.stride(1) and k.stride(0) == k.shape[1] * k.stride(1)
            and v.stride(0) == v.shape[1] * v.stride(1) and q.stride(1) % 8 == 0 and k.stride(1) % 8 == 0
            and q.data_ptr() % 16 == 0 and k.data_ptr() % 16 == 0 and v.data_ptr() % 16 == 0)


def _xattn_own_ok(q, k, v, d):
    return (not (torch.is_grad_enabled() and (q.requires_grad or k.requires_grad or v.requires_grad))
            and k.shape[1] <= _lib.lib().yat_xattn_max_keys() and d <= 128
            and q.stride(2) == 1 and k.stride(2) == 1 and v.stride(2) == 1 and k.stride(1) == v.stride(1)
            and q.stride(0) == q.shape[1] * q.stride(1) and k.stride(0) == k.shape[1] * k.stride(1)
            and v.stride(0) == v.shape[1] * v.stride(1) and q.stride(1) % 8 == 0 and k.stride(1) % 8 == 0
            and q.data_ptr() % 16 == 0 and k.data_ptr() % 16 == 0 and v.data_ptr() % 16 == 0)


class AttnProcessor2_0:
    """Softmax scaled-dot-product attention (diffusers ``AttnProcessor2_0`` semantics)."""
    fold_key_mask = True
    own_forward = True          # no-grad forwards (exploration trials, the frozen preservation pass, sampling): yat_xattn_fwd
    # training forward + backward on yat_xattn_fwd / yat_xattn_bwd instead of cuDNN (YAT_XATTN_OWN_BACKWARD=0 / 1 overrides)
    own_backward = os.environ.get("YAT_XATTN_OWN_BACKWARD", "1") == "1"

    def __call__(self, attn, hidden_states, encoder_hidden_states=None, attention_mask=None):
        B = hidden_states.shape[0]
        if encoder_hidden_states is None:
            encoder_hidden_states = hidden_states
        query = attn.to_q(hidden_states)
        key = attn.to_k(encoder_hidden_states)
        value = attn.to_v(encoder_hidden_states)
        d = query.shape[-1] // attn.heads
        if (attention_mask is not None and self.fold_key_mask and query.is_cuda and query.dtype == torch.bfloat16
                and attn.norm_q is None and attn.norm_k is None and d % 8 == 0
                and attention_mask.numel() == B * attention_mask.shape[-1]):
            if self.own_backward and torch.is_grad_enabled() and _xattn_layout_ok(query, key, value, d):
                out = _XAttnOwn.apply(query, key, value, _bias_f32(attention_mask, B, key.shape[1]), attn.heads)
            elif self.own_forward and _xattn_own_ok(query, key, value, d):
                out = _xattn_own_forward(query, key, value, attention_mask, attn.heads, d)
            else:
                out = _sdpa_key_bias_folded(query, key, value, attention_mask.view(B, 1, 1, -1).to(query.dtype),
                                            attn.heads, d)
            out = attn.to_out[0](out)
            return attn.to_out[1](out)
        query = query.view(B, -1, attn.heads, d).transpose(1, 2)
        key = key.view(B, -1, attn.heads, d).transpose(1, 2)
        value = value.view(B, -1, attn.heads, d).transpose(1, 2)
        if attn.norm_q is not None:
            query = attn.norm_q(query)
        if attn.norm_k is not None:
            key = attn.norm_k(key)
        mask = None
        if attention_mask is not None:
            # additive bias [B, 1, Lk] -> [B, 1, 1, Lk], broadcast over heads and queries
            mask = attention_mask.view(B, 1, -1, attention_mask.shape[-1]).to(query.dtype)
        out = F.scaled_dot_product_attention(query, key, value, attn_mask=mask, dropout_p=0.0, is_causal=False)
        out = out.transpose(1, 2).reshape(B, -1, attn.heads * d).to(query.dtype)
        out = attn.to_out[0](out)
        out = attn.to_out[1](out)
        return out


class Attention(nn.Module):
    def __init__(self, query_dim: int, cross_attention_dim: Optional[int] = None, heads: int = 8, dim_head: int = 64,
                 dropout: float = 0.0, bias: bool = False, out_bias: bool = True, qk_norm: Optional[str] = None,
                 kv_heads: Optional[int] = None, processor=None):
        super().__init__()
        if qk_norm is not None:
            raise ValueError("qk_norm is not supported in this build (SANA-1.6B: qk_norm=None)")
        self.inner_dim = dim_head * heads
        self.query_dim = query_dim
        self.cross_attention_dim = cross_attention_dim if cross_attention_dim is not None else query_dim
        self.heads = heads
        self.norm_q = None
        self.norm_k = None
        # nn.Linear parameters / keys; the bias gradient comes from yat_colsum_bf16 (linear.py)
        self.to_q = Linear(query_dim, self.inner_dim, bias=bias)
        self.to_k = Linear(self.cross_attention_dim, self.inner_dim, bias=bias)
        self.to_v = Linear(self.cross_attention_dim, self.inner_dim, bias=bias)
        self.to_out = nn.ModuleList([Linear(self.inner_dim, query_dim, bias=out_bias), nn.Dropout(dropout)])
        self.processor = processor if processor is not None else AttnProcessor2_0()

    def set_processor(self, processor) -> None:
        self.processor = processor

    def get_processor(self):
        return self.processor

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **kwargs):
        return self.processor(self, hidden_states, encoder_hidden_states=encoder_hidden_states,
                              attention_mask=attention_mask, **kwargs)
