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


class TinyDenoiser(torch.nn.Module):
    """A few-parameter denoiser with the SanaTransformer2DModel call signature (x_t, encoder_hidden_states, timestep,
    encoder_attention_mask) -> .sample, fp32 parameters, bf16 in / out.  The text enters through a MASKED mean, so the
    result does not depend on how far the captions are padded (512 in the reference, trimmed in this build)."""

    def __init__(self, channels, emb_dim, seed=0):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.mix = torch.nn.Parameter(torch.randn((channels, channels, 1, 1), generator=g) * channels ** -0.5)
        self.txt = torch.nn.Parameter(torch.randn((channels, emb_dim), generator=g) * emb_dim ** -0.5)
        self.tsc = torch.nn.Parameter(torch.randn((channels,), generator=g) * 0.1)

    def forward(self, x, encoder_hidden_states=None, timestep=None, encoder_attention_mask=None, **kw):
        import types
        m = encoder_attention_mask.to(torch.float32)
        if m.dim() == 3:                               # additive bias form [B, 1, L]: 0 on real tokens
            m = (m.squeeze(1) == 0).to(torch.float32)
        ctx = (encoder_hidden_states.float() * m.unsqueeze(-1)).sum(1) / m.sum(1, keepdim=True).clamp_min(1.0)
        h = torch.einsum("oc,bchw->bohw", self.mix[:, :, 0, 0], x.float())    # fp32 matmul (no TF32 conv path)
        h = h + torch.nn.functional.linear(ctx, self.txt)[:, :, None, None]
        h = h + (self.tsc[None] * (timestep.float() / 1000.0)[:, None])[:, :, None, None]
        return types.SimpleNamespace(sample=h.to(x.dtype))


def loop_batches(n_micro, batch_size, channels, hw, emb_dim, seed):
    """The micro-batches the step-loop golden feeds (fresh objects every call: the reference's CFG dropout writes into the
    embeddings list it is handed, common/trainer.py:321-323)."""
    import types
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_micro):
        lens = [int(v) for v in torch.randint(3, 40, (batch_size,), generator=g)]
        out.append(types.SimpleNamespace(
            ratio="1.0", repa_features=None,
            vae_features=torch.randn((batch_size, channels, hw, hw), generator=g).to(torch.bfloat16),
            embeddings=[(0.5 * torch.randn((L, emb_dim), generator=g)).to(torch.bfloat16) for L in lens]))
    return out


MODEL_CFG = dict(in_channels=4, out_channels=4, num_attention_heads=2, attention_head_dim=32, num_layers=2,
                 num_cross_attention_heads=2, cross_attention_head_dim=32, cross_attention_dim=64, caption_channels=16,
                 sample_size=8)


def stub_model(seed, cfg=MODEL_CFG):
    """The product's SANA model (fp32, CPU) with every block's attn1 / attn2 / ff replaced by the linear stand-ins of
    block_stubs.py and the op-by-op (torch) block path: what is left is the glue the forward owns."""
    from yat_b200.sana_transformer import SanaTransformer2DModel
    torch.manual_seed(seed)
    model = SanaTransformer2DModel(**cfg)
    C = cfg["num_attention_heads"] * cfg["attention_head_dim"]
    g = torch.Generator().manual_seed(seed + 1)
    w = lambda: torch.randn((C, C), generator=g) * C ** -0.5  # noqa: E731
    for blk in model.transformer_blocks:
        blk.attn1, blk.attn2, blk.ff = StubSelfAttn(w()), StubCrossAttn(w(), w()), StubFF(w())
        blk.block_ops = "torch"
    model.norm_out.block_ops = "torch"
    return model.eval()


class StubPixArtBlock(torch.nn.Module):
    """Stand-in for a PixArt block that takes the reference's keyword call (hidden_states, attention_mask=, encoder_hidden_states=,
    encoder_attention_mask=, timestep=, ...) and the product's positional one (x, enc, mask, timestep) alike."""

    def __init__(self, w):
        super().__init__()
        self.w = w

    def forward(self, hidden_states, *args, **kw):
        enc = kw["encoder_hidden_states"] if "encoder_hidden_states" in kw else args[0]
        temb = kw["timestep"] if "timestep" in kw else args[2]
        D = hidden_states.shape[-1]
        y = torch.tanh(torch.nn.functional.linear(hidden_states, self.w.to(hidden_states.dtype)))
        return y + 0.1 * enc.mean(dim=1, keepdim=True) + 0.1 * temb[:, None, :D]


PIXART_REPA_CFG = dict(num_attention_heads=2, attention_head_dim=32, in_channels=4, out_channels=8, num_layers=3,
                       cross_attention_dim=64, caption_channels=16, sample_size=8, patch_size=2, interpolation_scale=1,
                       repa_depth=1, repa_encoder_hidden_size=24, repa_projector_dim=40)


def stub_pixart_repa_model(seed, cfg=PIXART_REPA_CFG):
    from yat_b200.pixart_transformer import REPAPixArtTransformerModel
    torch.manual_seed(seed)
    model = REPAPixArtTransformerModel(**cfg)
    D = cfg["num_attention_heads"] * cfg["attention_head_dim"]
    g = torch.Generator().manual_seed(seed + 1)
    model.transformer_blocks = torch.nn.ModuleList(
        [StubPixArtBlock(torch.randn((D, D), generator=g) * D ** -0.5) for _ in range(cfg["num_layers"])])
    model.set_block_ops("torch")
    return model.eval()


def bucket_stream(seed=81, n=63):
    """The decoded shard elements both the reference loop and the product accumulator are fed (two shards)."""
    g = torch.Generator().manual_seed(seed)
    keys = ["1.0", "0.5", "2.0"]
    shapes = {"1.0": (4, 4), "0.5": (3, 6), "2.0": (6, 3)}
    out = []
    for i in range(n):
        k = keys[int(torch.randint(0, 3, (1,), generator=g))]
        L = int(torch.randint(1, 9, (1,), generator=g))
        out.append({"__key__": f"{i:07d}", "ratio": k, "latent.pt": torch.randn((2,) + shapes[k], generator=g).to(torch.bfloat16),
                    "emb.pt": torch.randn((L, 8), generator=g).to(torch.bfloat16)})
    return [out[: n // 2], out[n // 2:]]
