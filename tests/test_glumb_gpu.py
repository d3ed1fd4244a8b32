"""GPU parity of the fused GLUMBConv middle (yat_glumb_mid_*) against the reference's op chain
(SiLU -> depthwise 3x3 conv + bias -> chunk -> x * SiLU(gate); utils/patch_sana_attention_layers.py:106-111
-> diffusers GLUMBConv) evaluated by torch in fp64 on the same bf16 inputs.  Spec tolerance 2e-2
relative for bf16 activations / gradients; asserted tighter."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _reference(u, w, b, H, W):
    """u [B, H*W, 2Hd] (any float dtype) -> [B, H*W, Hd], NCHW torch ops."""
    B, N, C2 = u.shape
    x = u.view(B, H, W, C2).permute(0, 3, 1, 2)
    x = F.silu(x)
    x = F.conv2d(x, w.view(C2, 1, 3, 3), b, padding=1, groups=C2)
    xa, g = torch.chunk(x, 2, dim=1)
    return (xa * F.silu(g)).permute(0, 2, 3, 1).reshape(B, N, C2 // 2)


@pytest.mark.parametrize("B,H,W,Hd", [(2, 16, 16, 320), (1, 32, 32, 5600), (2, 22, 44, 64), (1, 1, 1, 8), (3, 5, 3, 72),
                                      (1, 2, 96, 16), (1, 3, 130, 8)])
def test_glumb_mid_forward_backward(cuda, B, H, W, Hd):
    from yat_b200.glumb import glumb_mid
    g = torch.Generator().manual_seed(B * 1000 + H * 10 + W)
    u = (torch.randn((B, H * W, 2 * Hd), generator=g) * 1.2).to(torch.bfloat16).to(cuda)
    w = (torch.randn((2 * Hd, 1, 3, 3), generator=g) * 0.3).to(torch.bfloat16).to(cuda)
    b = (torch.randn((2 * Hd,), generator=g) * 0.2).to(torch.bfloat16).to(cuda)
    dmid = torch.randn((B, H * W, Hd), generator=g).to(torch.bfloat16).to(cuda)

    ug, wg, bg = [t.clone().requires_grad_(True) for t in (u, w, b)]
    mid = glumb_mid(ug, wg, bg, H, W)
    mid.backward(dmid)

    ur, wr, br = [t.double().requires_grad_(True) for t in (u, w, b)]
    ref = _reference(ur, wr, br, H, W)
    ref.backward(dmid.double())
    assert mid.shape == (B, H * W, Hd) and mid.dtype == torch.bfloat16
    assert _rel(mid, ref) < 1e-2
    assert _rel(ug.grad, ur.grad) < 1.5e-2
    assert _rel(wg.grad, wr.grad) < 1.5e-2
    assert _rel(bg.grad, br.grad) < 1.5e-2
    assert wg.grad.shape == w.shape and bg.grad.shape == b.shape
    # deterministic
    ug2, wg2, bg2 = [t.clone().requires_grad_(True) for t in (u, w, b)]
    glumb_mid(ug2, wg2, bg2, H, W).backward(dmid)
    assert torch.equal(ug2.grad, ug.grad) and torch.equal(wg2.grad, wg.grad) and torch.equal(bg2.grad, bg.grad)


def test_glumb_mid_is_at_least_as_accurate_as_the_bf16_reference_chain(cuda):
    from yat_b200.glumb import glumb_mid
    B, H, W, Hd = 2, 32, 32, 640
    g = torch.Generator().manual_seed(7)
    u = (torch.randn((B, H * W, 2 * Hd), generator=g) * 1.2).to(torch.bfloat16).to(cuda)
    w = (torch.randn((2 * Hd, 1, 3, 3), generator=g) * 0.3).to(torch.bfloat16).to(cuda)
    b = (torch.randn((2 * Hd,), generator=g) * 0.2).to(torch.bfloat16).to(cuda)
    truth = _reference(u.double(), w.double(), b.double(), H, W)
    plain = _reference(u, w, b, H, W)                       # the reference's own bf16 op chain (cuDNN)
    got = glumb_mid(u, w, b, H, W)
    assert _rel(got, truth) <= 1.1 * _rel(plain, truth) + 1e-3


def test_glumbconv_module_token_path_matches_nchw_path(cuda):
    """GLUMBConv.forward_tokens (GEMM + fused middle + GEMM) vs GLUMBConv.forward (NCHW convs), same weights."""
    from yat_b200.sana_transformer import GLUMBConv
    torch.manual_seed(0)
    m = GLUMBConv(128, 128, 2.5).to(cuda, torch.bfloat16)
    B, H, W = 2, 16, 12
    x = torch.randn((B, H * W, 128), device=cuda).to(torch.bfloat16)
    dy = torch.randn((B, H * W, 128), device=cuda).to(torch.bfloat16)
    outs = []
    for path in ("tokens", "nchw", "nchw64"):
        mm = m if path != "nchw64" else GLUMBConv(128, 128, 2.5).to(cuda, torch.float64)
        if path == "nchw64":
            mm.load_state_dict({k: v.double() for k, v in m.state_dict().items()})
        mm.zero_grad(set_to_none=True)
        xi = x.to(next(mm.parameters()).dtype).clone().requires_grad_(True)
        if path == "tokens":
            y = mm.forward_tokens(xi, H, W)
        else:
            y = mm(xi.unflatten(1, (H, W)).permute(0, 3, 1, 2)).flatten(2, 3).permute(0, 2, 1)
        y.backward(dy.to(y.dtype))
        outs.append((y.detach(), xi.grad, mm.conv_inverted.weight.grad, mm.conv_depth.weight.grad,
                     mm.conv_depth.bias.grad, mm.conv_point.weight.grad))
    tok, nchw, truth = outs
    for a, p, t in zip(tok, nchw, truth):
        assert a.shape == t.shape
        assert _rel(a, t) <= 1.25 * _rel(p, t) + 2e-3, (_rel(a, t), _rel(p, t))


def test_glumb_argument_errors(cuda):
    from yat_b200.glumb import glumb_mid
    u = torch.zeros((1, 16, 32), dtype=torch.bfloat16, device=cuda)
    w = torch.zeros((32, 1, 3, 3), dtype=torch.bfloat16, device=cuda)
    b = torch.zeros((32,), dtype=torch.bfloat16, device=cuda)
    with pytest.raises(ValueError):
        glumb_mid(u, w, b, 4, 5)                   # H*W != N
    with pytest.raises(ValueError):
        glumb_mid(u.float(), w, b, 4, 4)
    with pytest.raises(ValueError):
        glumb_mid(u, w[:16], b, 4, 4)
    with pytest.raises(ValueError):
        glumb_mid(u.cpu(), w.cpu(), b.cpu(), 4, 4)
    with pytest.raises(ValueError):                 # Hd = 12 is not a multiple of 8
        glumb_mid(torch.zeros((1, 16, 24), dtype=torch.bfloat16, device=cuda),
                  torch.zeros((24, 1, 3, 3), dtype=torch.bfloat16, device=cuda),
                  torch.zeros((24,), dtype=torch.bfloat16, device=cuda), 4, 4)


@pytest.mark.parametrize("B,H,W,C,Hd", [(2, 16, 16, 128, 320), (1, 32, 32, 2240, 5600), (3, 5, 3, 40, 72)])
def test_inverted_glumb_mid_bias_gradient_from_the_backward_kernel(cuda, B, H, W, C, Hd):
    """conv_inverted + middle as one autograd node (inverted_glumb_mid): same forward and gradients as
    F.linear -> glumb_mid, with conv_inverted's bias gradient = column sum of du formed inside glumb_bwd_kernel
    (checked against the fp64 sum of the du the kernel wrote, and against the autograd chain)."""
    from yat_b200.glumb import glumb_mid, inverted_glumb_mid
    g = torch.Generator().manual_seed(B * 100 + C)
    x = torch.randn((B, H * W, C), generator=g).to(torch.bfloat16).to(cuda)
    wi = (torch.randn((2 * Hd, C), generator=g) * C ** -0.5).to(torch.bfloat16).to(cuda)
    bi = (torch.randn((2 * Hd,), generator=g) * 0.2).to(torch.bfloat16).to(cuda)
    w = (torch.randn((2 * Hd, 1, 3, 3), generator=g) * 0.3).to(torch.bfloat16).to(cuda)
    b = (torch.randn((2 * Hd,), generator=g) * 0.2).to(torch.bfloat16).to(cuda)
    dmid = torch.randn((B, H * W, Hd), generator=g).to(torch.bfloat16).to(cuda)

    a = [t.clone().requires_grad_(True) for t in (x, wi, bi, w, b)]
    mid = inverted_glumb_mid(*a, H, W)
    mid.backward(dmid)
    r = [t.clone().requires_grad_(True) for t in (x, wi, bi, w, b)]
    u = F.linear(r[0], r[1], r[2])
    u.retain_grad()
    ref = glumb_mid(u, r[3], r[4], H, W)
    ref.backward(dmid)
    assert torch.equal(mid, ref)
    assert torch.equal(a[3].grad, r[3].grad) and torch.equal(a[4].grad, r[4].grad)
    assert _rel(a[0].grad, r[0].grad) < 1e-3 and _rel(a[1].grad, r[1].grad) < 1e-3      # same GEMMs
    exact = u.grad.double().sum((0, 1))
    assert _rel(a[2].grad, exact) < 6e-3          # fp32 sum of the unrounded du, one bf16 rounding
    assert _rel(a[2].grad, exact) <= _rel(r[2].grad, exact) + 3e-3
    a2 = [t.clone().requires_grad_(True) for t in (x, wi, bi, w, b)]
    inverted_glumb_mid(*a2, H, W).backward(dmid)
    assert torch.equal(a2[2].grad, a[2].grad)     # deterministic
