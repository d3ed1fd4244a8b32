"""Deterministic stand-ins for the sub-modules of a SANA block (attn1, attn2, ff), shared by the reference-executed golden
generator and the tests: plain bf16 linear maps, so the block WRAPPER arithmetic (adaLN-single modulate, gate, residual,
utils/patch_sana_attention_layers.py:82-115) is what the comparison exercises."""
import torch


class StubSelfAttn(torch.nn.Module):
    def __init__(self, w):
        super().__init__()
        self.w = w

    def forward(self, x, **kw):
        return torch.nn.functional.linear(x, self.w.to(x.device))


class StubCrossAttn(torch.nn.Module):
    def __init__(self, w, we):
        super().__init__()
        self.w, self.we = w, we

    def forward(self, x, encoder_hidden_states=None, attention_mask=None, **kw):
        ctx = torch.nn.functional.linear(encoder_hidden_states.mean(dim=1, keepdim=True), self.we.to(x.device))
        return torch.nn.functional.linear(x, self.w.to(x.device)) + ctx


class StubFF(torch.nn.Module):
    """Per-pixel linear map: called NCHW by the reference block, token-major (forward_tokens) by the fused block."""

    def __init__(self, w):
        super().__init__()
        self.w = w

    def forward(self, x):                                   # [B, C, H, W]
        y = torch.nn.functional.linear(x.permute(0, 2, 3, 1), self.w.to(x.device))
        return y.permute(0, 3, 1, 2)

    def forward_tokens(self, x, height, width):             # [B, N, C]
        return torch.nn.functional.linear(x, self.w.to(x.device))
