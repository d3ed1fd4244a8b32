"""Functional restatement of SanaTransformer2DModel.forward (oracle; test infrastructure only).

Follows utils/patched_sana_transformer.py:88-167 (constructor / module inventory) and :229-349
(forward), the block body restated at utils/patch_sana_attention_layers.py:72-115, and the
diffusers building blocks those files import (PatchEmbed, AdaLayerNormSingle,
PixArtAlphaTextProjection, RMSNorm, Attention + SanaLinearAttnProcessor2_0 / AttnProcessor2_0,
GLUMBConv, SanaModulatedNorm) as recorded in SURVEY.md §8(c)(vi).  It is written as a pure
function of a HF-layout ``state_dict`` so that it shares no code with the product model
(yat_b200/sana_transformer.py); both must agree on the same weights.
"""
import math

import torch
import torch.nn.functional as F

from .linear_attention import sana_linear_attention_core


def timestep_embedding(timesteps, dim=256, max_period=10000):
    """diffusers get_timestep_embedding(flip_sin_to_cos=True, downscale_freq_shift=0, scale=1)."""
    half = dim // 2
    exponent = -math.log(max_period) * torch.arange(0, half, dtype=torch.float32, device=timesteps.device) / half
    emb = timesteps[:, None].float() * torch.exp(exponent)[None, :]
    return torch.cat([torch.cos(emb), torch.sin(emb)], dim=-1)


def rms_norm(x, weight, eps):
    dt = x.dtype
    var = x.to(torch.float32).pow(2).mean(-1, keepdim=True)
    x = x * torch.rsqrt(var + eps)
    if weight is not None:
        if weight.dtype in (torch.float16, torch.bfloat16):
            x = x.to(weight.dtype)
        x = x * weight
    else:
        x = x.to(dt)
    return x


def softmax_attention(x, enc, bias, sd, pre, heads):
    """diffusers AttnProcessor2_0 (cross attention, additive mask bias broadcast over heads/queries)."""
    B, L, _ = x.shape
    q = F.linear(x, sd[pre + "to_q.weight"], sd.get(pre + "to_q.bias"))
    k = F.linear(enc, sd[pre + "to_k.weight"], sd.get(pre + "to_k.bias"))
    v = F.linear(enc, sd[pre + "to_v.weight"], sd.get(pre + "to_v.bias"))
    d = q.shape[-1] // heads
    q = q.view(B, -1, heads, d).transpose(1, 2)
    k = k.view(B, -1, heads, d).transpose(1, 2)
    v = v.view(B, -1, heads, d).transpose(1, 2)
    mask = None if bias is None else bias[:, None, :, :].to(q.dtype)   # [B,1,1,Lk]
    o = F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=0.0, is_causal=False)
    o = o.transpose(1, 2).reshape(B, -1, heads * d).to(q.dtype)
    return F.linear(o, sd[pre + "to_out.0.weight"], sd.get(pre + "to_out.0.bias"))


def linear_attention(x, sd, pre, heads):
    q = F.linear(x, sd[pre + "to_q.weight"], sd.get(pre + "to_q.bias"))
    k = F.linear(x, sd[pre + "to_k.weight"], sd.get(pre + "to_k.bias"))
    v = F.linear(x, sd[pre + "to_v.weight"], sd.get(pre + "to_v.bias"))
    o = sana_linear_attention_core(q, k, v, heads)
    return F.linear(o, sd[pre + "to_out.0.weight"], sd.get(pre + "to_out.0.bias"))


def glumbconv(x, sd, pre):
    """diffusers GLUMBConv(norm_type=None, residual_connection=False) on NCHW input."""
    c2 = sd[pre + "conv_inverted.weight"].shape[0]
    x = F.conv2d(x, sd[pre + "conv_inverted.weight"], sd[pre + "conv_inverted.bias"])
    x = F.silu(x)
    x = F.conv2d(x, sd[pre + "conv_depth.weight"], sd[pre + "conv_depth.bias"], padding=1, groups=c2)
    x, gate = torch.chunk(x, 2, dim=1)
    x = x * F.silu(gate)
    return F.conv2d(x, sd[pre + "conv_point.weight"], None)


def sana_block(sd, pre, x, temb, enc, bias, heads, xheads, ph, pw, eps=1e-6, softmax_self_attention=False):
    """One SanaTransformerBlock, utils/patch_sana_attention_layers.py:82-115: x [B, N, C] tokens, temb [B, 6*C]
    (AdaLayerNormSingle output), enc [B, L, C] projected + normed text, bias [B, 1, L] additive key mask or None.
    ``pre`` = state-dict prefix ("transformer_blocks.<i>.")."""
    B, _, C = x.shape
    mod = (sd[pre + "scale_shift_table"][None] + temb.reshape(B, 6, -1)).chunk(6, dim=1)
    shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp = mod
    n = F.layer_norm(x, (C,), None, None, eps)
    n = (n * (1 + scale_msa) + shift_msa).to(x.dtype)
    if softmax_self_attention:
        a = softmax_attention(n, n, None, sd, pre + "attn1.", heads)
    else:
        a = linear_attention(n, sd, pre + "attn1.", heads)
    x = x + gate_msa * a
    a = softmax_attention(x, enc, bias, sd, pre + "attn2.", xheads)
    x = a + x
    n = F.layer_norm(x, (C,), None, None, eps)
    n = n * (1 + scale_mlp) + shift_mlp
    n = n.unflatten(1, (ph, pw)).permute(0, 3, 1, 2)
    f = glumbconv(n, sd, pre + "ff.")
    f = f.flatten(2, 3).permute(0, 2, 1)
    return x + gate_mlp * f


def sana_forward(sd, cfg, hidden_states, encoder_hidden_states, timestep, encoder_attention_mask=None,
                 gradient_checkpointing=False):
    """cfg: dict with num_attention_heads, attention_head_dim, num_layers, num_cross_attention_heads,
    patch_size, norm_eps, modified_blocks (softmax self-attention on those blocks,
    utils/patch_sana_attention_layers.py:125-137).  ``gradient_checkpointing`` recomputes each block
    in backward like the reference does (train_sana.py:63, patched:301-312).  Returns ``sample``."""
    heads, n_layers = cfg["num_attention_heads"], cfg["num_layers"]
    xheads, p = cfg["num_cross_attention_heads"], cfg.get("patch_size", 1)
    eps = cfg.get("norm_eps", 1e-6)
    modified = cfg.get("modified_blocks", [])
    dt = hidden_states.dtype

    bias = None
    if encoder_attention_mask is not None and encoder_attention_mask.ndim == 2:
        bias = ((1 - encoder_attention_mask.to(dt)) * -10000.0).unsqueeze(1)          # patched:275-277

    B, _, H, W = hidden_states.shape
    ph, pw = H // p, W // p
    x = F.conv2d(hidden_states, sd["patch_embed.proj.weight"], sd["patch_embed.proj.bias"], stride=p)
    x = x.flatten(2).transpose(1, 2)                                                  # patched:284

    t_proj = timestep_embedding(timestep).to(dt)
    e = F.linear(t_proj, sd["time_embed.emb.timestep_embedder.linear_1.weight"],
                 sd["time_embed.emb.timestep_embedder.linear_1.bias"])
    embedded_timestep = F.linear(F.silu(e), sd["time_embed.emb.timestep_embedder.linear_2.weight"],
                                 sd["time_embed.emb.timestep_embedder.linear_2.bias"])
    temb = F.linear(F.silu(embedded_timestep), sd["time_embed.linear.weight"], sd["time_embed.linear.bias"])

    enc = F.linear(encoder_hidden_states, sd["caption_projection.linear_1.weight"], sd["caption_projection.linear_1.bias"])
    enc = F.gelu(enc, approximate="tanh")
    enc = F.linear(enc, sd["caption_projection.linear_2.weight"], sd["caption_projection.linear_2.bias"])
    enc = enc.view(B, -1, x.shape[-1])
    enc = rms_norm(enc, sd["caption_norm.weight"], 1e-5)

    C = x.shape[-1]

    def block(i, x, temb, enc):
        return sana_block(sd, f"transformer_blocks.{i}.", x, temb, enc, bias, heads, xheads, ph, pw, eps,
                          softmax_self_attention=i in modified)

    for i in range(n_layers):
        if gradient_checkpointing and torch.is_grad_enabled():
            from torch.utils.checkpoint import checkpoint
            x = checkpoint(block, i, x, temb, enc, use_reentrant=False)
        else:
            x = block(i, x, temb, enc)

    shift, scale = (sd["scale_shift_table"][None] + embedded_timestep[:, None]).chunk(2, dim=1)
    x = F.layer_norm(x, (C,), None, None, 1e-6) * (1 + scale) + shift
    x = F.linear(x, sd["proj_out.weight"], sd["proj_out.bias"])
    x = x.reshape(B, ph, pw, p, p, -1).permute(0, 5, 1, 3, 2, 4)
    return x.reshape(B, -1, ph * p, pw * p)
