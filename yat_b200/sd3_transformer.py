"""SD3.5 MMDiT (diffusers ``SD3Transformer2DModel``) restated on the B200 block ops - SURVEY.md §8(f) N4.

The reference loads it through ``StableDiffusion3Pipeline.from_pretrained`` and trains ``pipe.transformer`` with the
flow-matching step of train_sd35.py:165-194 (``model(hidden_states, encoder_hidden_states, timestep,
pooled_projections)``).  diffusers is absent from this image, so the architecture is restated from its published
definition with the HF state-dict key layout: ``pos_embed`` (patch conv + a cropped window of a fixed 2-D sin-cos table),
``time_text_embed`` (timestep + pooled-text conditioning), ``context_embedder``, ``JointTransformerBlock`` x N
(``norm1`` = AdaLayerNormZero, or SD35AdaLayerNormZeroX with a second image-only attention ``attn2`` on the
``dual_attention_layers`` of SD3.5-medium; joint attention over [image ; text] tokens with RMS q/k norms; the last
block is ``context_pre_only``), ``norm_out`` (AdaLayerNormContinuous), ``proj_out``.  The modulated LayerNorms and the
gated residuals run as ``ln_modulate`` / ``gate_residual`` kernels, biased projections take their bias gradient from
``yat_colsum_bf16``; attention is torch SDPA (library).  ``oracle/sd3_model.py`` is the independent functional
restatement the tests compare with; parity with diffusers itself is unpinned (package absent).
"""
from types import SimpleNamespace

import torch
import torch.nn.functional as F
from torch import nn
from torch.utils.checkpoint import checkpoint

from .block_ops import gate_residual, ln_modulate
from .linear import Linear
from .pixart_transformer import FeedForward, sincos_2d
from .sana_transformer import RMSNorm, TimestepEmbedding, Timesteps, Transformer2DModelOutput


class PatchEmbedCropped(nn.Module):
    """Patch conv + the centre crop of a [max, max] sin-cos table (``pos_embed_max_size``): persistent buffer
    ``pos_embed`` like diffusers, so every aspect-ratio bucket reads the same table."""

    def __init__(self, sample_size, patch_size, in_channels, embed_dim, pos_embed_max_size):
        super().__init__()
        self.patch_size, self.max = patch_size, pos_embed_max_size
        self.proj = nn.Conv2d(in_channels, embed_dim, kernel_size=patch_size, stride=patch_size, bias=True)
        pe = sincos_2d(embed_dim, pos_embed_max_size, pos_embed_max_size, sample_size // patch_size, 1.0)
        self.register_buffer("pos_embed", pe.unsqueeze(0), persistent=True)

    def forward(self, latent):
        h, w = latent.shape[-2] // self.patch_size, latent.shape[-1] // self.patch_size
        if h > self.max or w > self.max:
            raise ValueError(f"latent grid {h}x{w} exceeds pos_embed_max_size {self.max}")
        top, left = (self.max - h) // 2, (self.max - w) // 2
        pe = self.pos_embed.reshape(1, self.max, self.max, -1)[:, top:top + h, left:left + w, :].reshape(1, h * w, -1)
        x = self.proj(latent).flatten(2).transpose(1, 2)
        return (x + pe).to(x.dtype)


class _TextProjection(nn.Module):          # PixArtAlphaTextProjection(act_fn="silu")
    def __init__(self, in_features, hidden_size):
        super().__init__()
        self.linear_1 = nn.Linear(in_features, hidden_size)
        self.act_1 = nn.SiLU()
        self.linear_2 = nn.Linear(hidden_size, hidden_size)

    def forward(self, x):
        return self.linear_2(self.act_1(self.linear_1(x)))


class CombinedTimestepTextProjEmbeddings(nn.Module):
    def __init__(self, embedding_dim, pooled_projection_dim):
        super().__init__()
        self.time_proj = Timesteps(256)
        self.timestep_embedder = TimestepEmbedding(256, embedding_dim)
        self.text_embedder = _TextProjection(pooled_projection_dim, embedding_dim)

    def forward(self, timestep, pooled_projection):
        t = self.timestep_embedder(self.time_proj(timestep).to(dtype=pooled_projection.dtype))
        return t + self.text_embedder(pooled_projection)


class _AdaNorm(nn.Module):
    """AdaLayerNormZero (6 chunks) / SD35AdaLayerNormZeroX (9 chunks) / AdaLayerNormContinuous (2 chunks, scale
    first): ``linear(silu(emb))`` -> [B, n, C]; the LayerNorm + modulate itself is applied by the caller's kernel."""

    def __init__(self, dim, chunks):
        super().__init__()
        self.chunks = chunks
        self.silu = nn.SiLU()
        self.linear = nn.Linear(dim, chunks * dim)
        self.norm = nn.LayerNorm(dim, elementwise_affine=False, eps=1e-6)

    def forward(self, emb):
        return self.linear(self.silu(emb)).view(emb.shape[0], self.chunks, -1)


def _modulate(norm, x, scale, shift, fused):
    if fused:
        return ln_modulate(x, scale, shift, norm.eps)
    return norm(x) * (1 + scale[:, None]) + shift[:, None]


def _gated_add(res, y, gate, fused):
    if fused:
        return gate_residual(res, y, gate)
    return res + gate[:, None] * y


class JointAttention(nn.Module):
    """diffusers ``Attention(added_kv_proj_dim=dim, qk_norm="rms_norm")`` + ``JointAttnProcessor2_0``."""

    def __init__(self, dim, heads, head_dim, joint=True, context_pre_only=False, qk_norm=True):
        super().__init__()
        inner = heads * head_dim
        self.heads, self.joint, self.context_pre_only = heads, joint, context_pre_only
        self.to_q, self.to_k, self.to_v = Linear(dim, inner), Linear(dim, inner), Linear(dim, inner)
        self.norm_q = RMSNorm(head_dim, eps=1e-6) if qk_norm else None
        self.norm_k = RMSNorm(head_dim, eps=1e-6) if qk_norm else None
        self.to_out = nn.ModuleList([Linear(inner, dim), nn.Dropout(0.0)])
        if joint:
            self.add_q_proj, self.add_k_proj, self.add_v_proj = Linear(dim, inner), Linear(dim, inner), Linear(dim, inner)
            self.norm_added_q = RMSNorm(head_dim, eps=1e-6) if qk_norm else None
            self.norm_added_k = RMSNorm(head_dim, eps=1e-6) if qk_norm else None
            if not context_pre_only:
                self.to_add_out = Linear(inner, dim)

    def _heads(self, x, norm):
        B, L, _ = x.shape
        x = x.view(B, L, self.heads, -1).transpose(1, 2)
        return norm(x) if norm is not None else x

    def forward(self, hidden_states, encoder_hidden_states=None):
        B, L, _ = hidden_states.shape
        q = self._heads(self.to_q(hidden_states), self.norm_q)
        k = self._heads(self.to_k(hidden_states), self.norm_k)
        v = self._heads(self.to_v(hidden_states), None)
        if encoder_hidden_states is not None:
            q = torch.cat([q, self._heads(self.add_q_proj(encoder_hidden_states), self.norm_added_q)], dim=2)
            k = torch.cat([k, self._heads(self.add_k_proj(encoder_hidden_states), self.norm_added_k)], dim=2)
            v = torch.cat([v, self._heads(self.add_v_proj(encoder_hidden_states), None)], dim=2)
        o = F.scaled_dot_product_attention(q, k, v, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(B, -1, q.shape[1] * q.shape[-1]).to(q.dtype)
        if encoder_hidden_states is None:
            return self.to_out[0](o)
        img, ctx = o[:, :L], o[:, L:]
        return self.to_out[0](img), (None if self.context_pre_only else self.to_add_out(ctx))


class JointTransformerBlock(nn.Module):
    def __init__(self, dim, heads, head_dim, context_pre_only=False, qk_norm=True, use_dual_attention=False):
        super().__init__()
        self.context_pre_only, self.use_dual_attention = context_pre_only, use_dual_attention
        self.norm1 = _AdaNorm(dim, 9 if use_dual_attention else 6)
        self.norm1_context = _AdaNorm(dim, 2 if context_pre_only else 6)
        self.attn = JointAttention(dim, heads, head_dim, joint=True, context_pre_only=context_pre_only, qk_norm=qk_norm)
        self.attn2 = JointAttention(dim, heads, head_dim, joint=False, qk_norm=qk_norm) if use_dual_attention else None
        self.norm2 = nn.LayerNorm(dim, elementwise_affine=False, eps=1e-6)
        self.ff = FeedForward(dim)
        if not context_pre_only:
            self.norm2_context = nn.LayerNorm(dim, elementwise_affine=False, eps=1e-6)
            self.ff_context = FeedForward(dim)
        self.block_ops = "sm100"

    def forward(self, hidden_states, encoder_hidden_states, temb):
        fused = self.block_ops == "sm100"
        m = self.norm1(temb)                                            # [B, 6 | 9, C]
        shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp = (m[:, i] for i in range(6))
        mc = self.norm1_context(temb)
        if self.context_pre_only:
            norm_enc = _modulate(self.norm1_context.norm, encoder_hidden_states, mc[:, 0], mc[:, 1], fused)   # scale, shift
        else:
            c_shift_msa, c_scale_msa, c_gate_msa, c_shift_mlp, c_scale_mlp, c_gate_mlp = (mc[:, i] for i in range(6))
            norm_enc = _modulate(self.norm1_context.norm, encoder_hidden_states, c_scale_msa, c_shift_msa, fused)
        if self.use_dual_attention and not fused:
            nh = self.norm1.norm(hidden_states)                         # one LayerNorm, two modulations
            norm_hidden = nh * (1 + scale_msa[:, None]) + shift_msa[:, None]
            norm_hidden2 = nh * (1 + m[:, 7][:, None]) + m[:, 6][:, None]
        else:
            norm_hidden = _modulate(self.norm1.norm, hidden_states, scale_msa, shift_msa, fused)
            norm_hidden2 = _modulate(self.norm1.norm, hidden_states, m[:, 7], m[:, 6], fused) if self.use_dual_attention else None
        attn_output, context_attn_output = self.attn(norm_hidden, norm_enc)
        hidden_states = _gated_add(hidden_states, attn_output, gate_msa, fused)
        if self.use_dual_attention:
            hidden_states = _gated_add(hidden_states, self.attn2(norm_hidden2), m[:, 8], fused)
        norm_hidden = _modulate(self.norm2, hidden_states, scale_mlp, shift_mlp, fused)
        hidden_states = _gated_add(hidden_states, self.ff(norm_hidden), gate_mlp, fused)
        if self.context_pre_only:
            return None, hidden_states
        encoder_hidden_states = _gated_add(encoder_hidden_states, context_attn_output, c_gate_msa, fused)
        norm_enc = _modulate(self.norm2_context, encoder_hidden_states, c_scale_mlp, c_shift_mlp, fused)
        encoder_hidden_states = _gated_add(encoder_hidden_states, self.ff_context(norm_enc), c_gate_mlp, fused)
        return encoder_hidden_states, hidden_states


SD3_CONFIGS = {
    # stabilityai/stable-diffusion-3.5-medium (MMDiT-X): BASELINE.json config 4
    "sd35_medium": dict(sample_size=128, patch_size=2, in_channels=16, num_layers=24, attention_head_dim=64,
                        num_attention_heads=24, joint_attention_dim=4096, caption_projection_dim=1536,
                        pooled_projection_dim=2048, out_channels=16, pos_embed_max_size=384,
                        dual_attention_layers=tuple(range(13)), qk_norm="rms_norm"),
    "sd3_tiny": dict(sample_size=16, patch_size=2, in_channels=16, num_layers=3, attention_head_dim=32,
                     num_attention_heads=2, joint_attention_dim=64, caption_projection_dim=64,
                     pooled_projection_dim=2048, out_channels=16, pos_embed_max_size=24,
                     dual_attention_layers=(0, 1), qk_norm="rms_norm"),
}


class SD3Transformer2DModel(nn.Module):
    def __init__(self, sample_size=128, patch_size=2, in_channels=16, num_layers=24, attention_head_dim=64,
                 num_attention_heads=24, joint_attention_dim=4096, caption_projection_dim=1536,
                 pooled_projection_dim=2048, out_channels=16, pos_embed_max_size=384, dual_attention_layers=(),
                 qk_norm=None):
        super().__init__()
        if qk_norm not in (None, "rms_norm"):
            raise ValueError(f"qk_norm {qk_norm!r} is not supported")
        inner = num_attention_heads * attention_head_dim
        if caption_projection_dim != inner:
            raise ValueError("caption_projection_dim must equal heads * head_dim")
        cfg = dict(locals())
        cfg.pop("self"), cfg.pop("__class__", None), cfg.pop("inner")
        cfg["dual_attention_layers"] = tuple(dual_attention_layers)
        cfg["caption_channels"] = joint_attention_dim
        self.config = SimpleNamespace(**cfg)
        self.pos_embed = PatchEmbedCropped(sample_size, patch_size, in_channels, inner, pos_embed_max_size)
        self.time_text_embed = CombinedTimestepTextProjEmbeddings(inner, pooled_projection_dim)
        self.context_embedder = nn.Linear(joint_attention_dim, caption_projection_dim)
        self.transformer_blocks = nn.ModuleList([
            JointTransformerBlock(inner, num_attention_heads, attention_head_dim, context_pre_only=i == num_layers - 1,
                                  qk_norm=qk_norm is not None, use_dual_attention=i in dual_attention_layers)
            for i in range(num_layers)])
        self.norm_out = _AdaNorm(inner, 2)
        self.proj_out = nn.Linear(inner, patch_size * patch_size * out_channels)
        self.gradient_checkpointing = False
        self.block_ops = "sm100"

    @property
    def dtype(self):
        return next(self.parameters()).dtype

    def enable_gradient_checkpointing(self):
        self.gradient_checkpointing = True

    def set_block_ops(self, impl):
        if impl not in ("sm100", "torch"):
            raise ValueError(f"unknown block_ops {impl!r}")
        for m in self.modules():
            if hasattr(m, "block_ops"):
                m.block_ops = impl

    def forward(self, hidden_states, encoder_hidden_states=None, timestep=None, pooled_projections=None,
                return_dict=True, **kw):
        B = hidden_states.shape[0]
        p = self.config.patch_size
        h, w = hidden_states.shape[-2] // p, hidden_states.shape[-1] // p
        x = self.pos_embed(hidden_states)
        temb = self.time_text_embed(timestep, pooled_projections.to(x.dtype))
        enc = self.context_embedder(encoder_hidden_states.to(x.dtype))
        use_ckpt = torch.is_grad_enabled() and self.gradient_checkpointing
        for block in self.transformer_blocks:
            if use_ckpt:
                enc, x = checkpoint(block, x, enc, temb, use_reentrant=False)
            else:
                enc, x = block(x, enc, temb)
        mo = self.norm_out(temb)
        x = _modulate(self.norm_out.norm, x, mo[:, 0], mo[:, 1], self.block_ops == "sm100")      # scale, shift
        x = self.proj_out(x)
        x = x.reshape(B, h, w, p, p, -1).permute(0, 5, 1, 3, 2, 4)
        out = x.reshape(B, -1, h * p, w * p)
        return Transformer2DModelOutput(sample=out) if return_dict else (out,)
