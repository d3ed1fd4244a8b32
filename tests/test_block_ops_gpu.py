"""GPU parity of the fused modulate / gate / residual kernels (yat_ln_modulate_*, yat_gate_residual_*)
against the reference's op-by-op expressions (utils/patch_sana_attention_layers.py:84-95,104-113)
evaluated in fp32/fp64 by torch.  Tolerance: bf16 outputs and gradients <= 2e-2 relative by spec
(asserted much tighter: the kernels compute in fp32 and round once)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-30)).item()


def _inputs(B, N, C, seed, dev):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn((B, N, C), generator=g) * 1.5 + 0.3).to(torch.bfloat16).to(dev)
    y = torch.randn((B, N, C), generator=g).to(torch.bfloat16).to(dev)
    mod = (torch.randn((B, 6, C), generator=g) * 0.5).to(torch.bfloat16).to(dev)   # [B, 6, C]: chunks are strided views
    dout = torch.randn((B, N, C), generator=g).to(torch.bfloat16).to(dev)
    return x, y, mod, dout


def _ref_ln_modulate(x, scale, shift, eps):
    xd = x.double()
    xh = torch.nn.functional.layer_norm(xd, (x.shape[-1],), eps=eps)
    return xh * (1 + scale.double()) + shift.double()


@pytest.mark.parametrize("B,N,C", [(2, 256, 128), (16, 1024, 2240), (3, 100, 2240), (1, 7, 64), (2, 33, 4096)])
def test_ln_modulate_forward_backward(cuda, B, N, C):
    from yat_b200.block_ops import ln_modulate
    x, _, mod, dout = _inputs(B, N, C, 1, cuda)
    shift, scale = mod.chunk(6, dim=1)[0], mod.chunk(6, dim=1)[1]
    xg = x.clone().requires_grad_(True)
    modg = mod.clone().requires_grad_(True)
    sh, sc = modg.chunk(6, dim=1)[0], modg.chunk(6, dim=1)[1]
    y = ln_modulate(xg, sc, sh, 1e-6)
    y.backward(dout)

    xr = x.double().requires_grad_(True)
    modr = mod.double().requires_grad_(True)
    yr = _ref_ln_modulate(xr, modr.chunk(6, dim=1)[1], modr.chunk(6, dim=1)[0], 1e-6)
    yr.backward(dout.double())
    assert _rel(y, yr) < 4e-3
    assert _rel(xg.grad, xr.grad) < 5e-3
    assert _rel(modg.grad[:, :2], modr.grad[:, :2]) < 5e-3
    assert torch.count_nonzero(modg.grad[:, 2:]) == 0
    # run-to-run identical (fixed reduction order)
    xg2 = x.clone().requires_grad_(True)
    modg2 = mod.clone().requires_grad_(True)
    ln_modulate(xg2, modg2.chunk(6, dim=1)[1], modg2.chunk(6, dim=1)[0], 1e-6).backward(dout)
    assert torch.equal(xg2.grad, xg.grad) and torch.equal(modg2.grad, modg.grad)
    del shift, scale


def test_ln_modulate_matches_reference_expression_in_bf16(cuda):
    """Against the reference's own bf16 op chain (LayerNorm -> *(1+scale) -> +shift, each rounding to bf16):
    the fused kernel rounds once, so the two agree to bf16 rounding noise."""
    from yat_b200.block_ops import ln_modulate
    x, _, mod, _ = _inputs(4, 512, 2240, 2, cuda)
    shift, scale = mod.chunk(6, dim=1)[0], mod.chunk(6, dim=1)[1]
    ref = torch.nn.functional.layer_norm(x, (2240,), eps=1e-6) * (1 + scale) + shift
    got = ln_modulate(x, scale, shift, 1e-6)
    assert _rel(got, ref) < 8e-3
    exact = _ref_ln_modulate(x, scale, shift, 1e-6)
    assert _rel(got, exact) <= _rel(ref, exact)   # and it is the more accurate of the two


@pytest.mark.parametrize("B,N,C", [(2, 256, 128), (16, 1024, 2240), (3, 100, 2240), (1, 7, 64)])
def test_gate_residual_forward_backward(cuda, B, N, C):
    from yat_b200.block_ops import gate_residual
    x, y, mod, dout = _inputs(B, N, C, 3, cuda)
    xg, yg = x.clone().requires_grad_(True), y.clone().requires_grad_(True)
    modg = mod.clone().requires_grad_(True)
    out = gate_residual(xg, yg, modg.chunk(6, dim=1)[2])
    out.backward(dout)
    xr, yr = x.double().requires_grad_(True), y.double().requires_grad_(True)
    modr = mod.double().requires_grad_(True)
    outr = xr + modr.chunk(6, dim=1)[2] * yr
    outr.backward(dout.double())
    assert _rel(out, outr) < 4e-3
    assert torch.equal(xg.grad, dout)                       # the residual branch passes the gradient through
    assert _rel(yg.grad, yr.grad) < 4e-3
    assert _rel(modg.grad[:, 2], modr.grad[:, 2]) < 5e-3
    assert torch.count_nonzero(modg.grad[:, :2]) == 0 and torch.count_nonzero(modg.grad[:, 3:]) == 0


def test_block_ops_argument_errors(cuda):
    from yat_b200.block_ops import gate_residual, ln_modulate
    x = torch.zeros((2, 8, 64), dtype=torch.bfloat16, device=cuda)
    m = torch.zeros((2, 1, 64), dtype=torch.bfloat16, device=cuda)
    with pytest.raises(ValueError):
        ln_modulate(x.float(), m, m)
    with pytest.raises(ValueError):
        ln_modulate(x, m[:, :, :32], m)
    with pytest.raises(ValueError):
        gate_residual(x, x[:, :4], m)
    with pytest.raises(ValueError):
        ln_modulate(x.cpu(), m.cpu(), m.cpu())
    with pytest.raises(ValueError):
        ln_modulate(torch.zeros((2, 8, 60), dtype=torch.bfloat16, device=cuda),
                    torch.zeros((2, 1, 60), dtype=torch.bfloat16, device=cuda),
                    torch.zeros((2, 1, 60), dtype=torch.bfloat16, device=cuda))


def test_fused_block_matches_reference_block(cuda):
    """A whole SanaTransformerBlock: fused block ops vs the reference's op-by-op expressions, same
    weights.  Truth = the op-by-op block in fp32; the fused bf16 block must be at least as close to
    it as the op-by-op bf16 block is (both within the 2e-2 spec on the output)."""
    from yat_b200.sana_transformer import SanaTransformerBlock
    torch.manual_seed(0)
    blk = SanaTransformerBlock(dim=128, num_attention_heads=4, attention_head_dim=32, num_cross_attention_heads=2,
                               cross_attention_head_dim=64, cross_attention_dim=128, attention_bias=False).to(cuda)
    B, H, W = 2, 16, 16
    x = torch.randn((B, H * W, 128), device=cuda).to(torch.bfloat16)
    enc = torch.randn((B, 40, 128), device=cuda).to(torch.bfloat16)
    temb = (torch.randn((B, 6 * 128), device=cuda) * 0.3).to(torch.bfloat16)

    def run(impl, dtype):
        b = SanaTransformerBlock(dim=128, num_attention_heads=4, attention_head_dim=32, num_cross_attention_heads=2,
                                 cross_attention_head_dim=64, cross_attention_dim=128, attention_bias=False)
        b.load_state_dict(blk.state_dict())
        b = b.to(cuda, dtype)
        b.block_ops = impl
        xi = x.to(dtype).requires_grad_(True)
        o = b(xi, None, enc.to(dtype), None, temb.to(dtype), H, W)
        o.float().square().mean().backward()
        return o.detach(), xi.grad, b.scale_shift_table.grad, b.attn1.to_q.weight.grad, b.ff.conv_point.weight.grad

    truth = run("torch", torch.float32)
    fused = run("sm100", torch.bfloat16)
    plain = run("torch", torch.bfloat16)
    assert _rel(fused[0], truth[0]) < 2e-2
    for f, p, t in zip(fused, plain, truth):
        assert _rel(f, t) <= 1.25 * _rel(p, t) + 2e-3, (_rel(f, t), _rel(p, t))


@pytest.mark.parametrize("rows,C", [(16384, 2240), (16 * 192, 2240), (1, 8), (33, 64), (100, 11200), (5000, 72)])
def test_colsum_and_linear_bias_gradient(cuda, rows, C):
    """yat_colsum_bf16 = grad_output.sum(0) of a biased Linear (fp32 accumulation, fixed order), and the Linear module
    built on it: same forward, same dx / dW GEMMs, bias gradient at least as close to fp64 as torch's reduce."""
    from yat_b200.linear import Linear, colsum
    g = torch.Generator().manual_seed(rows + C)
    a = torch.randn((rows, C), generator=g).to(torch.bfloat16).to(cuda)
    s = colsum(a)
    exact = a.double().sum(0)
    assert s.shape == (C,) and s.dtype == torch.bfloat16
    assert torch.equal(s, exact.to(torch.bfloat16)) or _rel(s, exact) < 4e-3
    assert torch.equal(colsum(a), s)
    if rows * C > 4_000_000 or C > 4096:
        return
    torch.manual_seed(0)
    lin = Linear(C, 64).to(cuda).to(torch.bfloat16)
    ref = torch.nn.Linear(C, 64).to(cuda).to(torch.bfloat16)
    ref.load_state_dict(lin.state_dict())
    x = a.view(1, rows, C).clone().requires_grad_(True)
    xr = a.view(1, rows, C).clone().requires_grad_(True)
    dy = torch.randn((1, rows, 64), generator=g).to(torch.bfloat16).to(cuda)
    y, yr = lin(x), ref(xr)
    assert torch.equal(y, yr)
    y.backward(dy)
    yr.backward(dy)
    assert _rel(x.grad, xr.grad) < 1e-3 and _rel(lin.weight.grad, ref.weight.grad) < 1e-3
    exact = dy.double().sum((0, 1))
    assert _rel(lin.bias.grad, exact) <= _rel(ref.bias.grad, exact) + 3e-3
    with pytest.raises(ValueError):
        colsum(a.cpu())


@pytest.mark.parametrize("B,N,C", [(2, 256, 128), (16, 1024, 2240), (3, 100, 2240), (2, 33, 4096)])
@pytest.mark.parametrize("with_add", [False, True])
def test_ln_modulate_res_folds_the_residual_arithmetic(cuda, B, N, C, with_add):
    """ln_modulate_res (yat_add_ln_modulate_fwd / yat_ln_modulate_bwd_res): h = x [+ x2] rounded to bf16 like the
    reference's tensor add (utils/patch_sana_attention_layers.py:103), y = ln_modulate(h); the gradients reaching h
    through the normalised branch and through the residual branch are combined inside the backward kernel.
    Checked against fp64 torch for the same graph; h is bit-identical to torch's bf16 add."""
    from yat_b200.block_ops import ln_modulate, ln_modulate_res
    x, x2, mod, dy = _inputs(B, N, C, 7, cuda)
    g = torch.Generator().manual_seed(9)
    dh = torch.randn((B, N, C), generator=g).to(torch.bfloat16).to(cuda)
    xg, ag = x.clone().requires_grad_(True), x2.clone().requires_grad_(True)
    modg = mod.clone().requires_grad_(True)
    sh, sc = modg.chunk(6, dim=1)[0], modg.chunk(6, dim=1)[1]
    y, h = ln_modulate_res(xg, sc, sh, 1e-6, add=ag if with_add else None)
    torch.autograd.backward([y, h], [dy, dh])
    h_ref_bf16 = (x + x2) if with_add else x
    assert torch.equal(h.detach(), h_ref_bf16)
    # the normalised branch equals the stand-alone op on the same (rounded) h, bit for bit
    assert torch.equal(y.detach(), ln_modulate(h_ref_bf16, mod.chunk(6, dim=1)[1], mod.chunk(6, dim=1)[0], 1e-6))
    # fp64 reference of the same graph (the rounding of h is part of the function: straight-through for the add)
    hr = h_ref_bf16.double().requires_grad_(True)
    modr = mod.double().requires_grad_(True)
    yr = _ref_ln_modulate(hr, modr.chunk(6, dim=1)[1], modr.chunk(6, dim=1)[0], 1e-6)
    torch.autograd.backward([yr, hr * 1.0], [dy.double(), dh.double()])
    assert _rel(xg.grad, hr.grad) < 5e-3
    if with_add:
        assert torch.equal(ag.grad, xg.grad)
    else:
        assert ag.grad is None
    assert _rel(modg.grad[:, :2], modr.grad[:, :2]) < 5e-3
    # one rounding instead of autograd's two (round dx_ln, then add dres in bf16): at least as close to fp64
    x3 = h_ref_bf16.clone().requires_grad_(True)
    y3 = ln_modulate(x3, mod.chunk(6, dim=1)[1], mod.chunk(6, dim=1)[0], 1e-6)
    torch.autograd.backward([y3, x3 * 1.0], [dy, dh])
    assert _rel(xg.grad, hr.grad) <= _rel(x3.grad, hr.grad) * 1.02
    # only one of the two outputs used downstream
    xg2 = x.clone().requires_grad_(True)
    y2, h2 = ln_modulate_res(xg2, mod.chunk(6, dim=1)[1], mod.chunk(6, dim=1)[0], 1e-6)
    y2.backward(dy)
    x4 = x.clone().requires_grad_(True)
    ln_modulate(x4, mod.chunk(6, dim=1)[1], mod.chunk(6, dim=1)[0], 1e-6).backward(dy)
    assert torch.equal(xg2.grad, x4.grad)
