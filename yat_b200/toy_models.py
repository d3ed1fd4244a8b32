"""Small signature-compatible stand-in denoisers for the PixArt-Sigma and SD3.5 hooks (TEST FIXTURES).

The real transformers are restated in ``pixart_transformer.py`` (PixArt-Sigma XL) and ``sd3_transformer.py``
(SD3.5-medium MMDiT-X) and are what the hooks build from ``params.model_config``.  These few-layer modules have the
same ``forward`` signatures and output shapes as ``PixArtTransformer2DModel`` (8 output channels = eps + learned
sigma, train_pixart_sigma.py:178-182) and ``SD3Transformer2DModel`` (``pooled_projections``, train_sd35.py:187-191):
the hook tests use them where the CPU oracle has to finish in seconds.
"""
from types import SimpleNamespace

import torch
from torch import nn


class _Out:
    def __init__(self, sample):
        self.sample = sample


class _TinyDenoiser(nn.Module):
    def __init__(self, in_channels, out_channels, caption_channels, hidden, pooled_channels=0):
        super().__init__()
        self.config = SimpleNamespace(in_channels=in_channels, out_channels=out_channels,
                                      caption_channels=caption_channels, sample_size=None)
        self.inp = nn.Conv2d(in_channels, hidden, 3, padding=1)
        self.text = nn.Linear(caption_channels, hidden)
        self.time = nn.Linear(1, hidden)
        self.pooled = nn.Linear(pooled_channels, hidden) if pooled_channels else None
        self.mid = nn.Conv2d(hidden, hidden, 3, padding=1)
        self.out = nn.Conv2d(hidden, out_channels, 3, padding=1)
        self.gradient_checkpointing = False

    @property
    def dtype(self):
        return next(self.parameters()).dtype

    def enable_gradient_checkpointing(self):
        self.gradient_checkpointing = True

    def _body(self, hidden_states, encoder_hidden_states, timestep, mask, pooled):
        dt = self.dtype
        h = self.inp(hidden_states.to(dt))
        enc = self.text(encoder_hidden_states.to(dt))
        if mask is not None:
            m = mask.to(dt).unsqueeze(-1)
            ctx = (enc * m).sum(1) / m.sum(1).clamp_min(1)
        else:
            ctx = enc.mean(1)
        ctx = ctx + self.time((timestep.to(torch.float32) / 1000.0).to(dt).unsqueeze(-1))
        if pooled is not None and self.pooled is not None:
            ctx = ctx + self.pooled(pooled.to(dt))
        h = torch.nn.functional.silu(h + ctx[:, :, None, None])
        h = torch.nn.functional.silu(self.mid(h))
        return _Out(self.out(h))


class ToyPixArtTransformer(_TinyDenoiser):
    """forward(hidden_states, encoder_hidden_states, timestep, encoder_attention_mask) -> .sample [B, 2C, h, w]."""

    def __init__(self, in_channels=4, caption_channels=4096, hidden=64):
        super().__init__(in_channels, 2 * in_channels, caption_channels, hidden)

    def forward(self, hidden_states, encoder_hidden_states=None, timestep=None, encoder_attention_mask=None, **kw):
        return self._body(hidden_states, encoder_hidden_states, timestep, encoder_attention_mask, None)


class ToySD3Transformer(_TinyDenoiser):
    """forward(hidden_states, encoder_hidden_states, timestep, pooled_projections) -> .sample [B, C, h, w]."""

    def __init__(self, in_channels=16, caption_channels=4096, pooled_channels=2048, hidden=64):
        super().__init__(in_channels, in_channels, caption_channels, hidden, pooled_channels)

    def forward(self, hidden_states, encoder_hidden_states=None, timestep=None, pooled_projections=None, **kw):
        return self._body(hidden_states, encoder_hidden_states, timestep, None, pooled_projections)
