"""Softmax cross-attention of the block's attn2 on the CPU (oracle; test infrastructure only).

Restates what diffusers' ``AttnProcessor2_0`` computes between the projections (reached from
utils/patch_sana_attention_layers.py:100-106 as ``self.attn2(hidden, encoder_hidden_states=..., attention_mask=...)``;
diffusers is absent from both boxes, so this follows its published algorithm - parity unpinned by the reference):

    heads split token-major [B, T, H*d] -> [B, H, T, d];  P = softmax(q k^T / sqrt(d) + bias);  out = P v  -> [B, Nq, H*d]

``bias`` is diffusers' additive mask ``(1 - mask) * -10000`` broadcast over heads and queries ([B, L]).  Everything in
``dtype`` (float64 for the truth the kernels are held to).  Also returns ln sum_j exp(.) per row (what the backward reads).
"""
import torch


def cross_attention(q, k, v, bias, heads, dtype=torch.float64, scale=None):
    B, Nq, C = q.shape
    L = k.shape[1]
    d = C // heads
    scale = d ** -0.5 if scale is None else scale
    qh = q.to(dtype).view(B, Nq, heads, d).transpose(1, 2)
    kh = k.to(dtype).view(B, L, heads, d).transpose(1, 2)
    vh = v.to(dtype).view(B, L, heads, d).transpose(1, 2)
    s = qh @ kh.transpose(-1, -2) * scale
    if bias is not None:
        s = s + bias.to(dtype).view(B, 1, 1, L)
    lse = torch.logsumexp(s, dim=-1)
    p = torch.softmax(s, dim=-1)
    out = (p @ vh).transpose(1, 2).reshape(B, Nq, C)
    return out, lse
