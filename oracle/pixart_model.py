"""Functional restatement of diffusers ``PixArtTransformer2DModel.forward`` (oracle; test infrastructure only).

The reference trains ``PixArtSigmaPipeline(...).transformer`` (train_pixart_sigma.py:26-40,151-185); diffusers is an
un-vendored dependency absent from this image, so this follows its published definition:
PatchEmbed with 2-D sin-cos positions (MAE order: the first half of the channels encodes the w coordinate),
AdaLayerNormSingle without the resolution / aspect-ratio conditions (PixArt-Sigma), PixArtAlphaTextProjection,
``BasicTransformerBlock(norm_type="ada_norm_single")`` x N (self-attention on the modulated norm1 with gate,
cross-attention on the un-normed stream without gate, feed-forward GELU-tanh x4 on the modulated norm2 with gate),
final LayerNorm + 2-row scale_shift_table, proj_out, un-patchify "nhwpqc->nchpwq".  Pure function of an HF-layout
state_dict: it shares no code with yat_b200/pixart_transformer.py.  Parity with diffusers: unpinned (package absent).
"""
import numpy as np
import torch
import torch.nn.functional as F

from .sana_model import softmax_attention, timestep_embedding


def sincos_pos_embed(dim, grid_h, grid_w, base_size, interpolation_scale):
    gh = np.arange(grid_h, dtype=np.float32) / (grid_h / base_size) / interpolation_scale
    gw = np.arange(grid_w, dtype=np.float32) / (grid_w / base_size) / interpolation_scale
    grid = np.stack(np.meshgrid(gw, gh), axis=0)                    # [2, H, W]: w coordinates first

    def one(d, pos):
        omega = np.arange(d // 2, dtype=np.float64)
        omega /= d / 2.0
        omega = 1.0 / 10000 ** omega
        out = np.einsum("m,d->md", pos.reshape(-1), omega)
        return np.concatenate([np.sin(out), np.cos(out)], axis=1)

    return torch.from_numpy(np.concatenate([one(dim // 2, grid[0]), one(dim // 2, grid[1])], axis=1)).float()


def pixart_forward(sd, cfg, hidden_states, encoder_hidden_states, timestep, encoder_attention_mask=None):
    heads, n_layers, p = cfg["num_attention_heads"], cfg["num_layers"], cfg["patch_size"]
    eps = cfg.get("norm_eps", 1e-6)
    dt = hidden_states.dtype
    bias = None
    if encoder_attention_mask is not None and encoder_attention_mask.ndim == 2:
        bias = ((1 - encoder_attention_mask.to(dt)) * -10000.0).unsqueeze(1)
    B, _, H, W = hidden_states.shape
    ph, pw = H // p, W // p
    x = F.conv2d(hidden_states, sd["pos_embed.proj.weight"], sd["pos_embed.proj.bias"], stride=p).flatten(2).transpose(1, 2)
    C = x.shape[-1]
    pe = sincos_pos_embed(C, ph, pw, cfg["sample_size"] // p, cfg["interpolation_scale"]).to(x.device)
    x = (x + pe[None]).to(dt)

    e = F.linear(timestep_embedding(timestep).to(dt), sd["adaln_single.emb.timestep_embedder.linear_1.weight"],
                 sd["adaln_single.emb.timestep_embedder.linear_1.bias"])
    embedded = F.linear(F.silu(e), sd["adaln_single.emb.timestep_embedder.linear_2.weight"],
                        sd["adaln_single.emb.timestep_embedder.linear_2.bias"])
    temb = F.linear(F.silu(embedded), sd["adaln_single.linear.weight"], sd["adaln_single.linear.bias"])
    enc = F.linear(encoder_hidden_states, sd["caption_projection.linear_1.weight"], sd["caption_projection.linear_1.bias"])
    enc = F.linear(F.gelu(enc, approximate="tanh"), sd["caption_projection.linear_2.weight"],
                   sd["caption_projection.linear_2.bias"]).view(B, -1, C)

    for i in range(n_layers):
        pre = f"transformer_blocks.{i}."
        shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp = (
            sd[pre + "scale_shift_table"][None] + temb.reshape(B, 6, -1)).chunk(6, dim=1)
        n = F.layer_norm(x, (C,), None, None, eps) * (1 + scale_msa) + shift_msa
        x = gate_msa * softmax_attention(n, n, None, sd, pre + "attn1.", heads) + x
        x = softmax_attention(x, enc, bias, sd, pre + "attn2.", heads) + x
        n = F.layer_norm(x, (C,), None, None, eps) * (1 + scale_mlp) + shift_mlp
        f = F.gelu(F.linear(n, sd[pre + "ff.net.0.proj.weight"], sd[pre + "ff.net.0.proj.bias"]), approximate="tanh")
        f = F.linear(f, sd[pre + "ff.net.2.weight"], sd[pre + "ff.net.2.bias"])
        x = gate_mlp * f + x

    shift, scale = (sd["scale_shift_table"][None] + embedded[:, None]).chunk(2, dim=1)
    x = F.layer_norm(x, (C,), None, None, 1e-6) * (1 + scale) + shift
    x = F.linear(x, sd["proj_out.weight"], sd["proj_out.bias"])
    x = x.reshape(B, ph, pw, p, p, -1).permute(0, 5, 1, 3, 2, 4)
    return x.reshape(B, -1, ph * p, pw * p)
