"""PixArt-Sigma transformer (diffusers ``PixArtTransformer2DModel``) restated on the B200 block ops - SURVEY.md §8(f) N4.

The reference loads it through ``PixArtSigmaPipeline.from_pretrained`` (train_pixart_sigma.py:26-40) and trains
``pipe.transformer`` with the DDPM step of train_pixart_sigma.py:151-185.  diffusers is absent from this image, so the
architecture is restated from its published definition (PixArt-Sigma XL-2-1024-MS: 28 ``BasicTransformerBlock``s with
``norm_type="ada_norm_single"``, 16 heads x 72, patch 2, sin-cos positions with ``interpolation_scale=2``, T5 captions
of 4096 channels, 8 output channels = eps + learned sigma) with the HF state-dict key layout (``pos_embed.proj``,
``adaln_single``, ``caption_projection``, ``transformer_blocks.N.{attn1,attn2,ff.net.0.proj,ff.net.2,scale_shift_table}``,
``scale_shift_table``, ``proj_out``), so a checkpoint of the reference's pipeline loads with ``load_state_dict``.
SANA's block is this block with linear attention and GLUMBConv swapped in, so the same kernels serve it:
``ln_modulate`` (norm1 / norm2 + adaLN-single modulate), ``gate_residual``, biased projections with the column-sum
bias gradient (``linear.Linear``); attention is torch SDPA (library).  ``oracle/pixart_model.py`` is the independent
functional restatement the tests compare with; parity with diffusers itself is unpinned (package absent).
"""
import math
from types import SimpleNamespace

import torch
import torch.nn.functional as F
from torch import nn
from torch.utils.checkpoint import checkpoint

from .attention import Attention, AttnProcessor2_0
from .block_ops import gate_residual, ln_modulate
from .linear import Linear
from .sana_transformer import AdaLayerNormSingle, PixArtAlphaTextProjection, Transformer2DModelOutput


def sincos_1d(dim, pos):
    """[M] positions -> [M, dim]: sin | cos halves, frequencies 10000^(-i / (dim/2)) (float64 like diffusers)."""
    omega = torch.arange(dim // 2, dtype=torch.float64) / (dim / 2.0)
    out = pos.reshape(-1).to(torch.float64)[:, None] * (1.0 / 10000 ** omega)[None, :]
    return torch.cat([torch.sin(out), torch.cos(out)], dim=1)


def sincos_2d(dim, grid_h, grid_w, base_size, interpolation_scale):
    """diffusers get_2d_sincos_pos_embed: token (h, w) -> [sincos(w coordinate) | sincos(h coordinate)], coordinates
    scaled by base_size / grid and 1 / interpolation_scale; returns [grid_h * grid_w, dim] float32, row-major in (h, w)."""
    gh = torch.arange(grid_h, dtype=torch.float32) / (grid_h / base_size) / interpolation_scale
    gw = torch.arange(grid_w, dtype=torch.float32) / (grid_w / base_size) / interpolation_scale
    ww, hh = torch.meshgrid(gw, gh, indexing="xy")               # both [grid_h, grid_w]
    return torch.cat([sincos_1d(dim // 2, ww), sincos_1d(dim // 2, hh)], dim=1).to(torch.float32)


class PatchEmbedSinCos(nn.Module):
    """Conv2d(k = s = patch) + flatten + fixed 2-D sin-cos positions (recomputed for other aspect-ratio buckets)."""

    def __init__(self, height, width, patch_size, in_channels, embed_dim, interpolation_scale):
        super().__init__()
        self.patch_size = patch_size
        self.height, self.width = height // patch_size, width // patch_size
        self.base_size = height // patch_size
        self.interpolation_scale = interpolation_scale
        self.embed_dim = embed_dim
        self.proj = nn.Conv2d(in_channels, embed_dim, kernel_size=patch_size, stride=patch_size, bias=True)
        pe = sincos_2d(embed_dim, self.height, self.width, self.base_size, interpolation_scale)
        self.register_buffer("pos_embed", pe.unsqueeze(0), persistent=False)
        self._cache = {}

    def forward(self, latent):
        h, w = latent.shape[-2] // self.patch_size, latent.shape[-1] // self.patch_size
        x = self.proj(latent).flatten(2).transpose(1, 2)
        if (h, w) == (self.height, self.width):
            pe = self.pos_embed
        else:
            key = (h, w, x.device)
            if key not in self._cache:
                self._cache[key] = sincos_2d(self.embed_dim, h, w, self.base_size,
                                             self.interpolation_scale).unsqueeze(0).to(x.device)
            pe = self._cache[key]
        return (x + pe).to(x.dtype)


class GELUProj(nn.Module):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = Linear(dim_in, dim_out)

    def forward(self, x):
        return F.gelu(self.proj(x), approximate="tanh")


class FeedForward(nn.Module):
    """diffusers FeedForward(activation_fn="gelu-approximate"): net = [GELU(proj), Dropout, Linear]."""

    def __init__(self, dim, mult=4):
        super().__init__()
        self.net = nn.ModuleList([GELUProj(dim, dim * mult), nn.Dropout(0.0), Linear(dim * mult, dim)])

    def forward(self, x):
        for m in self.net:
            x = m(x)
        return x


class PixArtTransformerBlock(nn.Module):
    """BasicTransformerBlock(norm_type="ada_norm_single"): modulated self-attention with gate, un-normed, un-gated
    cross-attention, modulated feed-forward with gate."""

    def __init__(self, dim, heads, head_dim, cross_attention_dim, attention_bias=True, norm_eps=1e-6):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, elementwise_affine=False, eps=norm_eps)
        self.attn1 = Attention(query_dim=dim, heads=heads, dim_head=head_dim, bias=attention_bias,
                               cross_attention_dim=None, processor=AttnProcessor2_0())
        self.norm2 = nn.LayerNorm(dim, elementwise_affine=False, eps=norm_eps)
        self.attn2 = Attention(query_dim=dim, cross_attention_dim=cross_attention_dim, heads=heads, dim_head=head_dim,
                               bias=attention_bias, processor=AttnProcessor2_0())
        self.ff = FeedForward(dim)
        self.scale_shift_table = nn.Parameter(torch.randn(6, dim) / dim ** 0.5)
        self.block_ops = "sm100"

    def forward(self, hidden_states, encoder_hidden_states, encoder_attention_mask, timestep):
        B = hidden_states.shape[0]
        shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp = (
            self.scale_shift_table[None] + timestep.reshape(B, 6, -1)).chunk(6, dim=1)
        fused = self.block_ops == "sm100"
        if fused:
            n = ln_modulate(hidden_states, scale_msa, shift_msa, self.norm1.eps)
            hidden_states = gate_residual(hidden_states, self.attn1(n), gate_msa)
        else:
            n = self.norm1(hidden_states) * (1 + scale_msa) + shift_msa
            hidden_states = gate_msa * self.attn1(n) + hidden_states
        hidden_states = self.attn2(hidden_states, encoder_hidden_states=encoder_hidden_states,
                                   attention_mask=encoder_attention_mask) + hidden_states
        if fused:
            n = ln_modulate(hidden_states, scale_mlp, shift_mlp, self.norm2.eps)
            return gate_residual(hidden_states, self.ff(n), gate_mlp)
        n = self.norm2(hidden_states) * (1 + scale_mlp) + shift_mlp
        return gate_mlp * self.ff(n) + hidden_states


PIXART_CONFIGS = {
    # PixArt-Sigma-XL-2-1024-MS (0.6 B): BASELINE.json config 3
    "pixart_sigma_xl_1024": dict(num_attention_heads=16, attention_head_dim=72, in_channels=4, out_channels=8,
                                 num_layers=28, cross_attention_dim=1152, caption_channels=4096, sample_size=128,
                                 patch_size=2, interpolation_scale=2),
    "pixart_tiny": dict(num_attention_heads=2, attention_head_dim=32, in_channels=4, out_channels=8, num_layers=2,
                        cross_attention_dim=64, caption_channels=64, sample_size=16, patch_size=2,
                        interpolation_scale=1),
}


class PixArtTransformer2DModel(nn.Module):
    def __init__(self, num_attention_heads=16, attention_head_dim=72, in_channels=4, out_channels=8, num_layers=28,
                 cross_attention_dim=1152, attention_bias=True, sample_size=128, patch_size=2, norm_eps=1e-6,
                 caption_channels=4096, interpolation_scale=None):
        super().__init__()
        inner = num_attention_heads * attention_head_dim
        if cross_attention_dim != inner:
            raise ValueError("PixArt projects captions to the model width: cross_attention_dim must equal heads * head_dim")
        if interpolation_scale is None:
            interpolation_scale = max(sample_size // 64, 1)
        cfg = dict(locals())
        cfg.pop("self"), cfg.pop("__class__", None), cfg.pop("inner")
        self.config = SimpleNamespace(**cfg)
        self.pos_embed = PatchEmbedSinCos(sample_size, sample_size, patch_size, in_channels, inner, interpolation_scale)
        self.transformer_blocks = nn.ModuleList([
            PixArtTransformerBlock(inner, num_attention_heads, attention_head_dim, cross_attention_dim, attention_bias,
                                   norm_eps) for _ in range(num_layers)])
        self.norm_out = nn.LayerNorm(inner, elementwise_affine=False, eps=1e-6)
        self.scale_shift_table = nn.Parameter(torch.randn(2, inner) / inner ** 0.5)
        self.proj_out = nn.Linear(inner, patch_size * patch_size * out_channels)
        self.adaln_single = AdaLayerNormSingle(inner)          # use_additional_conditions=False (Sigma)
        self.caption_projection = PixArtAlphaTextProjection(in_features=caption_channels, hidden_size=inner)
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

    def _after_block(self, index, hidden_states):
        """Tap for subclasses (REPAPixArtTransformerModel projects the tokens that leave one block)."""

    def forward(self, hidden_states, encoder_hidden_states=None, timestep=None, added_cond_kwargs=None,
                encoder_attention_mask=None, return_dict=True, **kw):
        if encoder_attention_mask is not None and encoder_attention_mask.ndim == 2:
            encoder_attention_mask = ((1 - encoder_attention_mask.to(hidden_states.dtype)) * -10000.0).unsqueeze(1)
        B = hidden_states.shape[0]
        p = self.config.patch_size
        h, w = hidden_states.shape[-2] // p, hidden_states.shape[-1] // p
        x = self.pos_embed(hidden_states)
        timestep, embedded_timestep = self.adaln_single(timestep, batch_size=B, hidden_dtype=x.dtype)
        enc = self.caption_projection(encoder_hidden_states).view(B, -1, x.shape[-1])
        use_ckpt = torch.is_grad_enabled() and self.gradient_checkpointing
        for index, block in enumerate(self.transformer_blocks):
            if use_ckpt:
                x = checkpoint(block, x, enc, encoder_attention_mask, timestep, use_reentrant=False)
            else:
                x = block(x, enc, encoder_attention_mask, timestep)
            self._after_block(index, x)
        shift, scale = (self.scale_shift_table[None] + embedded_timestep[:, None]).chunk(2, dim=1)
        if self.block_ops == "sm100":
            x = ln_modulate(x, scale, shift, self.norm_out.eps)
        else:
            x = self.norm_out(x) * (1 + scale) + shift
        x = self.proj_out(x)
        x = x.reshape(B, h, w, p, p, -1).permute(0, 5, 1, 3, 2, 4)          # "nhwpqc->nchpwq"
        out = x.reshape(B, -1, h * p, w * p)
        return Transformer2DModelOutput(sample=out) if return_dict else (out,)


class ThreeLayerMLP(nn.Module):
    """REPA projector of the PixArt variant (utils/patch_pixart_sigma_transformer.py:10-22): Linear-SiLU-Linear-SiLU-Linear,
    parameters under ``fc.0 / fc.2 / fc.4``."""

    def __init__(self, input_size, projector_size, output_size):
        super().__init__()
        self.fc = nn.Sequential(nn.Linear(input_size, projector_size), nn.SiLU(),
                                nn.Linear(projector_size, projector_size), nn.SiLU(),
                                nn.Linear(projector_size, output_size))

    def forward(self, x):
        return self.fc(x)


class REPAPixArtTransformerModel(PixArtTransformer2DModel):
    """PixArt-Sigma with a REPA projection head (utils/patch_pixart_sigma_transformer.py:24-198; arXiv 2410.06940): the
    tokens that leave block ``repa_depth`` (0-based: the reference compares its block counter AFTER the block ran, :165)
    go through ``repa_mlps`` and stay on ``self.repa_proj`` [B, T, repa_encoder_hidden_size] until the next forward.
    Upstream nothing reads ``repa_proj`` (REPA is unfinished there); here ``calculate_loss`` turns it into the
    cosine alignment with the teacher tokens (``yat_repa_cos``, the same term as ``yat_b200.repa.RepaModel``)."""

    def __init__(self, *args, repa_depth=4, repa_encoder_hidden_size=768, repa_projector_dim=2048, **kwargs):
        super().__init__(*args, **kwargs)
        self.repa_depth = repa_depth
        self.repa_projector_dim = repa_projector_dim
        self.repa_encoder_hidden_size = repa_encoder_hidden_size
        self.config.repa_depth, self.config.repa_projector_dim = repa_depth, repa_projector_dim
        self.config.repa_encoder_hidden_size = repa_encoder_hidden_size
        self.repa_mlps = ThreeLayerMLP(self.config.cross_attention_dim, repa_projector_dim, repa_encoder_hidden_size)
        self.repa_proj = None

    def _after_block(self, index, hidden_states):
        if index == self.repa_depth:
            N, T, D = hidden_states.shape
            self.repa_proj = self.repa_mlps(hidden_states.reshape(-1, D)).reshape(N, T, -1)

    def calculate_loss(self, features, grid_hw, teacher_hw=(16, 16)):
        """-mean cos(avgpool(repa_proj), features); clears the stash (common/repa.py:150-152 does the same)."""
        from .repa import repa_cosine_loss
        if self.repa_proj is None:
            raise RuntimeError("calculate_loss needs a forward pass first (repa_proj is empty)")
        loss = repa_cosine_loss(self.repa_proj.to(torch.bfloat16), features, grid_hw, teacher_hw)
        self.repa_proj = None
        return loss
