"""Functional restatement of diffusers ``SD3Transformer2DModel.forward`` (oracle; test infrastructure only).

The reference trains ``StableDiffusion3Pipeline(...).transformer`` (train_sd35.py:165-194); diffusers is an un-vendored
dependency absent from this image, so this follows its published definition: PatchEmbed with the
centre crop of a ``pos_embed_max_size`` sin-cos table, CombinedTimestepTextProjEmbeddings, context_embedder,
JointTransformerBlock (AdaLayerNormZero / SD35AdaLayerNormZeroX / AdaLayerNormContinuous, JointAttnProcessor2_0 with
RMS q/k norms, image-only ``attn2`` on the dual-attention layers, ``context_pre_only`` last block), norm_out, proj_out.
Pure function of an HF-layout state_dict: shares no code with yat_b200/sd3_transformer.py.  Parity with diffusers:
unpinned (package absent).
"""
import torch
import torch.nn.functional as F

from .pixart_model import sincos_pos_embed
from .sana_model import rms_norm, timestep_embedding


def _lin(sd, name, x):
    return F.linear(x, sd[name + ".weight"], sd.get(name + ".bias"))


def _ln(x):
    return F.layer_norm(x, (x.shape[-1],), None, None, 1e-6)


def _ff(sd, pre, x):
    return _lin(sd, pre + "net.2", F.gelu(_lin(sd, pre + "net.0.proj", x), approximate="tanh"))


def _heads(x, heads, w):
    B, L, _ = x.shape
    x = x.view(B, L, heads, -1).transpose(1, 2)
    return rms_norm(x, w, 1e-6) if w is not None else x


def _attention(sd, pre, heads, x, ctx, context_pre_only):
    L = x.shape[1]
    q = _heads(_lin(sd, pre + "to_q", x), heads, sd.get(pre + "norm_q.weight"))
    k = _heads(_lin(sd, pre + "to_k", x), heads, sd.get(pre + "norm_k.weight"))
    v = _heads(_lin(sd, pre + "to_v", x), heads, None)
    if ctx is not None:
        q = torch.cat([q, _heads(_lin(sd, pre + "add_q_proj", ctx), heads, sd.get(pre + "norm_added_q.weight"))], dim=2)
        k = torch.cat([k, _heads(_lin(sd, pre + "add_k_proj", ctx), heads, sd.get(pre + "norm_added_k.weight"))], dim=2)
        v = torch.cat([v, _heads(_lin(sd, pre + "add_v_proj", ctx), heads, None)], dim=2)
    o = F.scaled_dot_product_attention(q, k, v)
    o = o.transpose(1, 2).reshape(x.shape[0], -1, q.shape[1] * q.shape[-1])
    if ctx is None:
        return _lin(sd, pre + "to_out.0", o)
    return _lin(sd, pre + "to_out.0", o[:, :L]), (None if context_pre_only else _lin(sd, pre + "to_add_out", o[:, L:]))


def sd3_forward(sd, cfg, hidden_states, encoder_hidden_states, timestep, pooled_projections):
    heads, n_layers, p = cfg["num_attention_heads"], cfg["num_layers"], cfg["patch_size"]
    dual = set(cfg.get("dual_attention_layers", ()))
    dt = hidden_states.dtype
    B, _, H, W = hidden_states.shape
    ph, pw = H // p, W // p
    x = F.conv2d(hidden_states, sd["pos_embed.proj.weight"], sd["pos_embed.proj.bias"], stride=p).flatten(2).transpose(1, 2)
    C, mx = x.shape[-1], cfg["pos_embed_max_size"]
    table = sincos_pos_embed(C, mx, mx, cfg["sample_size"] // p, 1.0).view(mx, mx, C)
    assert torch.allclose(table.reshape(1, mx * mx, C), sd["pos_embed.pos_embed"].float().cpu(), atol=1e-2)
    top, left = (mx - ph) // 2, (mx - pw) // 2
    x = (x + table[top:top + ph, left:left + pw].reshape(1, ph * pw, C).to(x.device)).to(dt)

    t = _lin(sd, "time_text_embed.timestep_embedder.linear_1", timestep_embedding(timestep).to(dt))
    t = _lin(sd, "time_text_embed.timestep_embedder.linear_2", F.silu(t))
    pz = _lin(sd, "time_text_embed.text_embedder.linear_2", F.silu(_lin(sd, "time_text_embed.text_embedder.linear_1", pooled_projections)))
    temb = t + pz
    enc = _lin(sd, "context_embedder", encoder_hidden_states)

    def ada(name, n):
        return _lin(sd, name + ".linear", F.silu(temb)).chunk(n, dim=1)

    for i in range(n_layers):
        pre = f"transformer_blocks.{i}."
        last = i == n_layers - 1
        if i in dual:
            s_msa, sc_msa, g_msa, s_mlp, sc_mlp, g_mlp, s_msa2, sc_msa2, g_msa2 = ada(pre + "norm1", 9)
        else:
            s_msa, sc_msa, g_msa, s_mlp, sc_mlp, g_mlp = ada(pre + "norm1", 6)
        nx = _ln(x)
        n1 = nx * (1 + sc_msa[:, None]) + s_msa[:, None]
        if last:
            c_scale, c_shift = ada(pre + "norm1_context", 2)
            nc = _ln(enc) * (1 + c_scale)[:, None] + c_shift[:, None]
        else:
            cs_msa, csc_msa, cg_msa, cs_mlp, csc_mlp, cg_mlp = ada(pre + "norm1_context", 6)
            nc = _ln(enc) * (1 + csc_msa[:, None]) + cs_msa[:, None]
        a, ca = _attention(sd, pre + "attn.", heads, n1, nc, last)
        x = x + g_msa.unsqueeze(1) * a
        if i in dual:
            n2 = nx * (1 + sc_msa2[:, None]) + s_msa2[:, None]
            x = x + g_msa2.unsqueeze(1) * _attention(sd, pre + "attn2.", heads, n2, None, False)
        x = x + g_mlp.unsqueeze(1) * _ff(sd, pre + "ff.", _ln(x) * (1 + sc_mlp[:, None]) + s_mlp[:, None])
        if last:
            enc = None
        else:
            enc = enc + cg_msa.unsqueeze(1) * ca
            enc = enc + cg_mlp.unsqueeze(1) * _ff(sd, pre + "ff_context.", _ln(enc) * (1 + csc_mlp[:, None]) + cs_mlp[:, None])

    scale, shift = ada("norm_out", 2)
    x = _ln(x) * (1 + scale)[:, None] + shift[:, None]
    x = _lin(sd, "proj_out", x)
    x = x.reshape(B, ph, pw, p, p, -1).permute(0, 5, 1, 3, 2, 4)
    return x.reshape(B, -1, ph * p, pw * p)
