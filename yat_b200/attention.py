"""Attention module and processors (diffusers attention-processor protocol).

The reference installs processors with ``attn.set_processor(...)`` and calls them as
``processor(attn, hidden_states, encoder_hidden_states=None, attention_mask=None)``
(utils/patch_sana_attention_layers.py:125-137).  ``Attention`` below carries the attribute names
processors and PEFT configs rely on (``to_q to_k to_v to_out.0``, ``heads``, ``norm_q``,
``norm_k``); ``SanaLinearAttnProcessorB200`` is the slot the tcgen05 kernel drops into, and
``AttnProcessor2_0`` (softmax SDPA, library call) is what ``patch_sana_attention_layers`` swaps in
for ``modified_blocks`` and what cross-attention uses.
"""
from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from .linear import Linear
from .linear_attention import sana_linear_attention


class SanaLinearAttnProcessorB200:
    """Drop-in for diffusers ``SanaLinearAttnProcessor2_0``: projections stay ``attn.to_q/k/v``
    (possibly PEFT-wrapped), the ReLU linear-attention core runs as one tcgen05 kernel."""

    def __init__(self, impl: str = "sm100"):
        self.impl = impl

    def __call__(self, attn, hidden_states, encoder_hidden_states=None, attention_mask=None):
        original_dtype = hidden_states.dtype
        if encoder_hidden_states is None:
            encoder_hidden_states = hidden_states
        if attn.norm_q is not None or attn.norm_k is not None:
            raise ValueError("qk_norm is not supported by the B200 linear-attention kernel "
                             "(SANA-1.6B uses qk_norm=None, utils/patched_sana_transformer.py:109)")
        query = attn.to_q(hidden_states)
        key = attn.to_k(encoder_hidden_states)
        value = attn.to_v(encoder_hidden_states)
        out = sana_linear_attention(query.to(torch.bfloat16), key.to(torch.bfloat16), value.to(torch.bfloat16),
                                    attn.heads, impl=self.impl)
        out = out.to(original_dtype)
        out = attn.to_out[0](out)
        out = attn.to_out[1](out)
        return out


class AttnProcessor2_0:
    """Softmax scaled-dot-product attention (diffusers ``AttnProcessor2_0`` semantics)."""

    def __call__(self, attn, hidden_states, encoder_hidden_states=None, attention_mask=None):
        B = hidden_states.shape[0]
        if encoder_hidden_states is None:
            encoder_hidden_states = hidden_states
        query = attn.to_q(hidden_states)
        key = attn.to_k(encoder_hidden_states)
        value = attn.to_v(encoder_hidden_states)
        d = query.shape[-1] // attn.heads
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
