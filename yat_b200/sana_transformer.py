"""SANA transformer with B200 linear attention.

Host-side mirror of utils/patched_sana_transformer.py (``PatchedSanaTransformer2DModel``: ctor /
config keys :88-112, ``attn_processors`` / ``set_attn_processor`` :169-227, ``forward`` signature
and body :229-349) and utils/patch_sana_attention_layers.py (``patch_sana_attention_layers``
:125-137, ``unfreeze_sana_blocks`` :117-123).  Module names match the Hugging Face checkpoints
(``patch_embed.proj``, ``time_embed.emb.timestep_embedder.linear_{1,2}``, ``time_embed.linear``,
``caption_projection.linear_{1,2}``, ``caption_norm``, ``transformer_blocks.i.{attn1,attn2}.*``,
``transformer_blocks.i.ff.{conv_inverted,conv_depth,conv_point}``, ``scale_shift_table``,
``proj_out``) so SANA-1.6B weights load unchanged.  GEMMs / SDPA / convolutions are library calls
(out of the hot-path scope); self-attention on unpatched blocks runs the
tcgen05 kernel through ``SanaLinearAttnProcessorB200``.
"""
import json
import math
import os
from types import SimpleNamespace
from typing import Any, Dict, Optional, Tuple, Union

import torch
import torch.nn.functional as F
from torch import nn
from torch.utils.checkpoint import checkpoint

from .attention import Attention, AttnProcessor2_0, SanaLinearAttnProcessorB200
from .block_ops import gate_residual, ln_modulate, ln_modulate_res
from .adapters import adapter_term
from .glumb import glumb_mid, inverted_glumb_mid


class Transformer2DModelOutput:
    def __init__(self, sample):
        self.sample = sample

    def __getitem__(self, i):
        return (self.sample,)[i]


class PatchEmbed(nn.Module):
    """Conv2d(k=p, s=p) + flatten; no positional embedding (interpolation_scale=None)."""

    def __init__(self, patch_size=1, in_channels=32, embed_dim=2240):
        super().__init__()
        self.patch_size = patch_size
        self.proj = nn.Conv2d(in_channels, embed_dim, kernel_size=patch_size, stride=patch_size, bias=True)

    def forward(self, latent):
        return self.proj(latent).flatten(2).transpose(1, 2)


class Timesteps(nn.Module):
    def __init__(self, num_channels=256):
        super().__init__()
        self.num_channels = num_channels

    def forward(self, timesteps):
        half = self.num_channels // 2
        exponent = -math.log(10000) * torch.arange(0, half, dtype=torch.float32, device=timesteps.device) / half
        emb = timesteps[:, None].float() * torch.exp(exponent)[None, :]
        return torch.cat([torch.cos(emb), torch.sin(emb)], dim=-1)   # flip_sin_to_cos=True


class TimestepEmbedding(nn.Module):
    def __init__(self, in_channels, time_embed_dim):
        super().__init__()
        self.linear_1 = nn.Linear(in_channels, time_embed_dim)
        self.act = nn.SiLU()
        self.linear_2 = nn.Linear(time_embed_dim, time_embed_dim)

    def forward(self, sample):
        return self.linear_2(self.act(self.linear_1(sample)))


class _TimestepSizeEmbeddings(nn.Module):
    def __init__(self, embedding_dim):
        super().__init__()
        self.time_proj = Timesteps(256)
        self.timestep_embedder = TimestepEmbedding(256, embedding_dim)

    def forward(self, timestep, hidden_dtype):
        return self.timestep_embedder(self.time_proj(timestep).to(dtype=hidden_dtype))


class AdaLayerNormSingle(nn.Module):
    def __init__(self, embedding_dim):
        super().__init__()
        self.emb = _TimestepSizeEmbeddings(embedding_dim)
        self.silu = nn.SiLU()
        self.linear = nn.Linear(embedding_dim, 6 * embedding_dim, bias=True)

    def forward(self, timestep, batch_size=None, hidden_dtype=None):
        embedded_timestep = self.emb(timestep, hidden_dtype)
        return self.linear(self.silu(embedded_timestep)), embedded_timestep


class PixArtAlphaTextProjection(nn.Module):
    def __init__(self, in_features, hidden_size):
        super().__init__()
        self.linear_1 = nn.Linear(in_features, hidden_size, bias=True)
        self.act_1 = nn.GELU(approximate="tanh")
        self.linear_2 = nn.Linear(hidden_size, hidden_size, bias=True)

    def forward(self, caption):
        return self.linear_2(self.act_1(self.linear_1(caption)))


class RMSNorm(nn.Module):
    def __init__(self, dim, eps, elementwise_affine=True):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim)) if elementwise_affine else None

    def forward(self, hidden_states):
        input_dtype = hidden_states.dtype
        variance = hidden_states.to(torch.float32).pow(2).mean(-1, keepdim=True)
        hidden_states = hidden_states * torch.rsqrt(variance + self.eps)
        if self.weight is not None:
            if self.weight.dtype in (torch.float16, torch.bfloat16):
                hidden_states = hidden_states.to(self.weight.dtype)
            hidden_states = hidden_states * self.weight
        else:
            hidden_states = hidden_states.to(input_dtype)
        return hidden_states


class GLUMBConv(nn.Module):
    """1x1 expand -> SiLU -> depthwise 3x3 -> GLU -> 1x1 project (no norm, no residual)."""

    def __init__(self, in_channels, out_channels, expand_ratio=2.5):
        super().__init__()
        hidden = int(expand_ratio * in_channels)
        self.nonlinearity = nn.SiLU()
        self.conv_inverted = nn.Conv2d(in_channels, hidden * 2, 1, 1, 0)
        self.conv_depth = nn.Conv2d(hidden * 2, hidden * 2, 3, 1, 1, groups=hidden * 2)
        self.conv_point = nn.Conv2d(hidden, out_channels, 1, 1, 0, bias=False)

    def forward(self, hidden_states):
        hidden_states = self.conv_inverted(hidden_states)
        hidden_states = self.nonlinearity(hidden_states)
        hidden_states = self.conv_depth(hidden_states)
        hidden_states, gate = torch.chunk(hidden_states, 2, dim=1)
        hidden_states = hidden_states * self.nonlinearity(gate)
        return self.conv_point(hidden_states)

    def forward_tokens(self, hidden_states, height, width):
        """Same module on token-major data [B, H*W, C] (the product path): the 1x1 convolutions are
        GEMMs on the conv weights viewed as [out, in], the middle is one fused kernel."""
        w_inv, w_pt = self.conv_inverted.weight, self.conv_point.weight
        t = adapter_term(self.conv_inverted, hidden_states)                # LoRA / LoKr term, None without an adapter
        if t is None and self.conv_inverted.bias is not None:
            mid = inverted_glumb_mid(hidden_states, w_inv.view(w_inv.shape[0], w_inv.shape[1]), self.conv_inverted.bias,
                                     self.conv_depth.weight, self.conv_depth.bias, height, width)
        else:
            u = F.linear(hidden_states, w_inv.view(w_inv.shape[0], w_inv.shape[1]), self.conv_inverted.bias)
            if t is not None:
                u = u + t
            mid = glumb_mid(u, self.conv_depth.weight, self.conv_depth.bias, height, width)
        out = F.linear(mid, w_pt.view(w_pt.shape[0], w_pt.shape[1]), self.conv_point.bias)
        t = adapter_term(self.conv_point, mid)
        return out if t is None else out + t


class SanaModulatedNorm(nn.Module):
    def __init__(self, dim, elementwise_affine=False, eps=1e-6):
        super().__init__()
        self.norm = nn.LayerNorm(dim, elementwise_affine=elementwise_affine, eps=eps)
        self.block_ops = "sm100"

    def forward(self, hidden_states, temb, scale_shift_table):
        shift, scale = (scale_shift_table[None] + temb[:, None].to(scale_shift_table.device)).chunk(2, dim=1)
        if self.block_ops == "sm100" and self.norm.weight is None:
            return ln_modulate(hidden_states, scale, shift, self.norm.eps)   # one kernel, fp32 math
        hidden_states = self.norm(hidden_states)
        return hidden_states * (1 + scale) + shift


class SanaTransformerBlock(nn.Module):
    """Block body as restated at utils/patch_sana_attention_layers.py:72-115."""

    def __init__(self, dim=2240, num_attention_heads=70, attention_head_dim=32, dropout=0.0,
                 num_cross_attention_heads=20, cross_attention_head_dim=112, cross_attention_dim=2240,
                 attention_bias=True, norm_elementwise_affine=False, norm_eps=1e-6, attention_out_bias=True,
                 mlp_ratio=2.5, qk_norm=None):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim, elementwise_affine=False, eps=norm_eps)
        self.attn1 = Attention(query_dim=dim, heads=num_attention_heads, dim_head=attention_head_dim,
                               dropout=dropout, bias=attention_bias, cross_attention_dim=None, qk_norm=qk_norm,
                               processor=SanaLinearAttnProcessorB200())
        self.attn2 = None
        if cross_attention_dim is not None:
            self.norm2 = nn.LayerNorm(dim, elementwise_affine=norm_elementwise_affine, eps=norm_eps)
            self.attn2 = Attention(query_dim=dim, cross_attention_dim=cross_attention_dim,
                                   heads=num_cross_attention_heads, dim_head=cross_attention_head_dim,
                                   dropout=dropout, bias=True, out_bias=attention_out_bias, qk_norm=qk_norm,
                                   processor=AttnProcessor2_0())
        self.ff = GLUMBConv(dim, dim, mlp_ratio)
        self.scale_shift_table = nn.Parameter(torch.randn(6, dim) / dim ** 0.5)
        # "sm100": fused CUDA kernels (the product path, CUDA bf16 only, raises otherwise);
        # "torch": the reference's op-by-op expressions (CPU structure tests)
        self.block_ops = "sm100" if not norm_elementwise_affine else "torch"

    def forward(self, hidden_states, attention_mask=None, encoder_hidden_states=None, encoder_attention_mask=None,
                timestep=None, height=None, width=None):
        batch_size = hidden_states.shape[0]
        shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp = (
            self.scale_shift_table[None] + timestep.reshape(batch_size, 6, -1)).chunk(6, dim=1)

        if self.block_ops == "sm100":
            return self._forward_fused(hidden_states, encoder_hidden_states, encoder_attention_mask, height, width,
                                       shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp)
        norm_hidden_states = self.norm1(hidden_states)
        norm_hidden_states = norm_hidden_states * (1 + scale_msa) + shift_msa
        norm_hidden_states = norm_hidden_states.to(hidden_states.dtype)
        attn_output = self.attn1(norm_hidden_states)
        hidden_states = hidden_states + gate_msa * attn_output

        if self.attn2 is not None:
            attn_output = self.attn2(hidden_states, encoder_hidden_states=encoder_hidden_states,
                                     attention_mask=encoder_attention_mask)
            hidden_states = attn_output + hidden_states

        norm_hidden_states = self.norm2(hidden_states)
        norm_hidden_states = norm_hidden_states * (1 + scale_mlp) + shift_mlp
        norm_hidden_states = norm_hidden_states.unflatten(1, (height, width)).permute(0, 3, 1, 2)
        ff_output = self.ff(norm_hidden_states)
        ff_output = ff_output.flatten(2, 3).permute(0, 2, 1)
        return hidden_states + gate_mlp * ff_output

    def _forward_fused(self, hidden_states, encoder_hidden_states, encoder_attention_mask, height, width,
                       shift_msa, scale_msa, gate_msa, shift_mlp, scale_mlp, gate_mlp):
        """Same block with the modulate / gate / residual arithmetic as fused kernels
        (yat_b200/csrc/block_ops.cu); everything stays token-major [B, N, C]."""
        # ln_modulate_res returns the residual-branch tensor from the same autograd node as the normalised one: the
        # two gradients of every residual-stream tensor are then added inside the LN backward kernel, and the attn2
        # residual add (:103) is formed inside the norm2 forward kernel - no torch elementwise pass is left here
        norm_hidden_states, hidden_states = ln_modulate_res(hidden_states, scale_msa, shift_msa, self.norm1.eps)
        attn_output = self.attn1(norm_hidden_states)
        hidden_states = gate_residual(hidden_states, attn_output, gate_msa)
        attn_output = None
        if self.attn2 is not None:
            attn_output = self.attn2(hidden_states, encoder_hidden_states=encoder_hidden_states,
                                     attention_mask=encoder_attention_mask)
        norm_hidden_states, hidden_states = ln_modulate_res(hidden_states, scale_mlp, shift_mlp, self.norm2.eps,
                                                            add=attn_output)
        ff_output = self.ff.forward_tokens(norm_hidden_states, height, width)
        return gate_residual(hidden_states, ff_output, gate_mlp)


def unfreeze_sana_blocks(transformer, layers):
    """utils/patch_sana_attention_layers.py:117-123."""
    for param in transformer.parameters():
        param.requires_grad = False
    for idx in layers:
        for param in transformer.transformer_blocks[idx].parameters():
            param.requires_grad = True


def patch_sana_attention_layers(transformer, layers):
    """utils/patch_sana_attention_layers.py:125-137: the listed blocks trade linear attention for
    softmax ``AttnProcessor2_0`` on attn1 and attn2; the list is recorded in the config."""
    for idx in layers:
        block = transformer.transformer_blocks[idx]
        block.attn1.set_processor(AttnProcessor2_0())
        block.attn2.set_processor(AttnProcessor2_0())
        transformer.transformer_blocks[idx] = block
    modified_blocks = list(transformer.config.modified_blocks)
    for l in layers:
        if l not in modified_blocks:
            modified_blocks.append(l)
    transformer.register_to_config(modified_blocks=modified_blocks)
    return transformer


class PatchedSanaTransformer2DModel(nn.Module):
    _supports_gradient_checkpointing = True
    _no_split_modules = ["SanaTransformerBlock", "PatchEmbed", "SanaModulatedNorm"]
    config_name = "config.json"
    weights_name = "diffusion_pytorch_model.safetensors"

    def __init__(self, in_channels: int = 32, out_channels: Optional[int] = 32, num_attention_heads: int = 70,
                 attention_head_dim: int = 32, num_layers: int = 20, num_cross_attention_heads: Optional[int] = 20,
                 cross_attention_head_dim: Optional[int] = 112, cross_attention_dim: Optional[int] = 2240,
                 caption_channels: int = 2304, mlp_ratio: float = 2.5, dropout: float = 0.0,
                 attention_bias: bool = False, sample_size: int = 32, patch_size: int = 1,
                 norm_elementwise_affine: bool = False, norm_eps: float = 1e-6,
                 interpolation_scale: Optional[int] = None, guidance_embeds: bool = False,
                 guidance_embeds_scale: float = 0.1, qk_norm: Optional[str] = None, timestep_scale: float = 1.0,
                 modified_blocks=()) -> None:
        super().__init__()
        if guidance_embeds:
            raise ValueError("guidance_embeds (SANA-Sprint) is outside this build's scope")
        if interpolation_scale is not None:
            raise ValueError("sincos positional embedding (interpolation_scale) is outside this build's scope")
        cfg = dict(locals())
        cfg.pop("self"), cfg.pop("__class__", None)
        cfg["modified_blocks"] = []
        self.config = SimpleNamespace(**cfg)

        out_channels = out_channels or in_channels
        inner_dim = num_attention_heads * attention_head_dim
        self.patch_embed = PatchEmbed(patch_size=patch_size, in_channels=in_channels, embed_dim=inner_dim)
        self.time_embed = AdaLayerNormSingle(inner_dim)
        self.caption_projection = PixArtAlphaTextProjection(in_features=caption_channels, hidden_size=inner_dim)
        self.caption_norm = RMSNorm(inner_dim, eps=1e-5, elementwise_affine=True)
        self.transformer_blocks = nn.ModuleList([
            SanaTransformerBlock(inner_dim, num_attention_heads, attention_head_dim, dropout=dropout,
                                 num_cross_attention_heads=num_cross_attention_heads,
                                 cross_attention_head_dim=cross_attention_head_dim,
                                 cross_attention_dim=cross_attention_dim, attention_bias=attention_bias,
                                 norm_elementwise_affine=norm_elementwise_affine, norm_eps=norm_eps,
                                 mlp_ratio=mlp_ratio, qk_norm=qk_norm)
            for _ in range(num_layers)])
        patch_sana_attention_layers(self, list(modified_blocks))
        self.scale_shift_table = nn.Parameter(torch.randn(2, inner_dim) / inner_dim ** 0.5)
        self.norm_out = SanaModulatedNorm(inner_dim, elementwise_affine=False, eps=1e-6)
        self.proj_out = nn.Linear(inner_dim, patch_size * patch_size * out_channels)
        self.gradient_checkpointing = False

    # ---- config / persistence (diffusers-style surface) ----
    def register_to_config(self, **kwargs):
        for k, v in kwargs.items():
            setattr(self.config, k, v)

    @property
    def dtype(self):
        return next(self.parameters()).dtype

    @property
    def device(self):
        return next(self.parameters()).device

    def enable_gradient_checkpointing(self):
        self.gradient_checkpointing = True

    def disable_gradient_checkpointing(self):
        self.gradient_checkpointing = False

    def save_pretrained(self, save_directory):
        from safetensors.torch import save_file
        os.makedirs(save_directory, exist_ok=True)
        cfg = dict(vars(self.config))
        cfg["_class_name"] = type(self).__name__
        with open(os.path.join(save_directory, self.config_name), "w") as f:
            json.dump(cfg, f, indent=2)
        save_file({k: v.detach().contiguous().cpu() for k, v in self.state_dict().items()},
                  os.path.join(save_directory, self.weights_name))

    @classmethod
    def from_pretrained(cls, path, torch_dtype=None, device_map=None, **kwargs):
        from safetensors.torch import load_file
        with open(os.path.join(path, cls.config_name)) as f:
            cfg = {k: v for k, v in json.load(f).items() if not k.startswith("_")}
        model = cls(**cfg)
        model.load_state_dict(load_file(os.path.join(path, cls.weights_name)))
        if torch_dtype is not None:
            model = model.to(torch_dtype)
        if device_map is not None:
            model = model.to(device_map)
        return model

    def set_block_ops(self, impl: str):
        """"sm100" (fused kernels, default) or "torch" (op-by-op reference expressions)."""
        if impl not in ("sm100", "torch"):
            raise ValueError(f"unknown block_ops {impl!r}")
        for m in self.modules():
            if hasattr(m, "block_ops"):
                m.block_ops = impl

    # ---- attention processors (utils/patched_sana_transformer.py:169-227) ----
    @property
    def attn_processors(self) -> Dict[str, Any]:
        processors = {}

        def add(name, module):
            if hasattr(module, "get_processor"):
                processors[f"{name}.processor"] = module.get_processor()
            for sub_name, child in module.named_children():
                add(f"{name}.{sub_name}", child)

        for name, module in self.named_children():
            add(name, module)
        return processors

    def set_attn_processor(self, processor):
        count = len(self.attn_processors.keys())
        if isinstance(processor, dict) and len(processor) != count:
            raise ValueError(
                f"A dict of processors was passed, but the number of processors {len(processor)} does not match the"
                f" number of attention layers: {count}. Please make sure to pass {count} processor classes.")

        def assign(name, module):
            if hasattr(module, "set_processor"):
                module.set_processor(processor if not isinstance(processor, dict) else processor.pop(f"{name}.processor"))
            for sub_name, child in module.named_children():
                assign(f"{name}.{sub_name}", child)

        for name, module in self.named_children():
            assign(name, module)

    # ---- forward (utils/patched_sana_transformer.py:229-349) ----
    def forward(self, hidden_states: torch.Tensor, encoder_hidden_states: torch.Tensor, timestep: torch.Tensor,
                guidance: Optional[torch.Tensor] = None, encoder_attention_mask: Optional[torch.Tensor] = None,
                attention_mask: Optional[torch.Tensor] = None, attention_kwargs: Optional[Dict[str, Any]] = None,
                controlnet_block_samples: Optional[Tuple[torch.Tensor]] = None,
                return_dict: bool = True) -> Union[Tuple[torch.Tensor, ...], Transformer2DModelOutput]:
        if guidance is not None:
            raise ValueError("guidance embeddings are outside this build's scope")
        if attention_mask is not None and attention_mask.ndim == 2:
            attention_mask = (1 - attention_mask.to(hidden_states.dtype)) * -10000.0
            attention_mask = attention_mask.unsqueeze(1)
        if encoder_attention_mask is not None and encoder_attention_mask.ndim == 2:
            encoder_attention_mask = (1 - encoder_attention_mask.to(hidden_states.dtype)) * -10000.0
            encoder_attention_mask = encoder_attention_mask.unsqueeze(1)

        batch_size, num_channels, height, width = hidden_states.shape
        p = self.config.patch_size
        post_patch_height, post_patch_width = height // p, width // p

        hidden_states = self.patch_embed(hidden_states)
        timestep, embedded_timestep = self.time_embed(timestep, batch_size=batch_size,
                                                      hidden_dtype=hidden_states.dtype)
        encoder_hidden_states = self.caption_projection(encoder_hidden_states)
        encoder_hidden_states = encoder_hidden_states.view(batch_size, -1, hidden_states.shape[-1])
        encoder_hidden_states = self.caption_norm(encoder_hidden_states)

        use_ckpt = torch.is_grad_enabled() and self.gradient_checkpointing
        for index_block, block in enumerate(self.transformer_blocks):
            if use_ckpt:
                hidden_states = checkpoint(block, hidden_states, attention_mask, encoder_hidden_states,
                                           encoder_attention_mask, timestep, post_patch_height, post_patch_width,
                                           use_reentrant=False)
            else:
                hidden_states = block(hidden_states, attention_mask, encoder_hidden_states, encoder_attention_mask,
                                      timestep, post_patch_height, post_patch_width)
            if controlnet_block_samples is not None and 0 < index_block <= len(controlnet_block_samples):
                hidden_states = hidden_states + controlnet_block_samples[index_block - 1]

        hidden_states = self.norm_out(hidden_states, embedded_timestep, self.scale_shift_table)
        hidden_states = self.proj_out(hidden_states)
        hidden_states = hidden_states.reshape(batch_size, post_patch_height, post_patch_width, p, p, -1)
        hidden_states = hidden_states.permute(0, 5, 1, 3, 2, 4)
        output = hidden_states.reshape(batch_size, -1, post_patch_height * p, post_patch_width * p)
        if not return_dict:
            return (output,)
        return Transformer2DModelOutput(sample=output)


SanaTransformer2DModel = PatchedSanaTransformer2DModel
