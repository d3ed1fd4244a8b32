"""GPU: fused denoising loss (yat_loss_fused) vs the CPU oracle, forward value and dL/dpred."""
import pytest
import torch

from oracle import losses, philox

pytestmark = pytest.mark.gpu


def _mk(shape, seed, scale=1.0):
    return (scale * torch.randn(shape, generator=torch.Generator().manual_seed(seed))).to(torch.bfloat16)


def _oracle(pred, x0, noise, **kw):
    p = pred.clone().requires_grad_(True)
    total, l_noise, l_pres = losses.denoise_loss(p, x0, noise, **kw)
    total.float().backward()
    return total.detach(), l_noise.detach(), l_pres.detach(), p.grad


def _rel(a, b):
    return ((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-30)).item()


@pytest.mark.parametrize("shape", [(16, 32, 32, 32), (2, 32, 16, 16), (8, 16, 128, 128), (3, 4, 88, 176)])
def test_fp32_velocity_mse_matches_reference(cuda, shape):
    """train_sana.py:217-218: MSELoss(pred.float(), (noise - latents).float()), <= 1e-5 relative."""
    from yat_b200.loss import fused_denoise_loss
    pred, x0, noise = _mk(shape, 1), _mk(shape, 2), _mk(shape, 3)
    total, _, _, g_ref = _oracle(pred, x0, noise, mode="velocity")
    p = pred.to(cuda).requires_grad_(True)
    loss, terms = fused_denoise_loss(p, x0.to(cuda), noise.to(cuda), return_terms=True)
    loss.backward()
    assert loss.dtype == torch.float32
    assert abs(loss.item() - total.item()) <= 1e-5 * abs(total.item())
    assert terms[2].item() == 0.0
    # bf16 gradient: identical rounding chain up to fp32 op order -> norm-wise << 2e-2, elementwise 1 bf16 ulp
    assert _rel(p.grad.cpu(), g_ref) < 2e-3
    assert ((p.grad.cpu().float() - g_ref.float()).abs() <= 2.0 ** -7 * g_ref.float().abs() + 1e-30).all()


def test_fp32_loss_is_deterministic(cuda):
    from yat_b200.loss import fused_denoise_loss
    shape = (16, 32, 32, 32)
    pred, x0, noise = _mk(shape, 1).to(cuda), _mk(shape, 2).to(cuda), _mk(shape, 3).to(cuda)
    a = fused_denoise_loss(pred, x0, noise)
    for _ in range(5):
        assert torch.equal(a, fused_denoise_loss(pred, x0, noise))


def test_bf16_loss_sd35_and_pixart(cuda):
    """train_sd35.py:193 (bf16 velocity MSE) and train_pixart_sigma.py:182-184 (eps MSE on the first
    half of an 8-channel prediction); <= 2e-2 relative, dpred of the unused half is exactly zero."""
    from yat_b200.loss import EPS, VELOCITY, fused_denoise_loss
    shape = (8, 16, 64, 64)
    pred, x0, noise = _mk(shape, 11), _mk(shape, 12), _mk(shape, 13)
    total, _, _, g_ref = _oracle(pred, x0, noise, mode="velocity", lowp=True)
    p = pred.to(cuda).requires_grad_(True)
    loss = fused_denoise_loss(p, x0.to(cuda), noise.to(cuda), target_mode=VELOCITY, lowp=True)
    loss.float().backward()
    assert loss.dtype == torch.bfloat16
    assert abs(loss.float().item() - total.float().item()) <= 2e-2 * abs(total.float().item())
    assert _rel(p.grad.cpu(), g_ref) < 2e-2

    lat = _mk((4, 4, 128, 128), 21)
    nz = _mk((4, 4, 128, 128), 22)
    pred8 = _mk((4, 8, 128, 128), 23)
    q = pred8.clone().requires_grad_(True)
    npred, tgt = losses.pixart_crop(q, nz)
    ref = losses.mse_lowp(npred, tgt)
    ref.float().backward()
    p8 = pred8.to(cuda).requires_grad_(True)
    loss8 = fused_denoise_loss(p8, None, nz.to(cuda), target_mode=EPS, lowp=True)
    loss8.float().backward()
    assert abs(loss8.float().item() - ref.float().item()) <= 2e-2 * abs(ref.float().item())
    assert (p8.grad[:, 4:] == 0).all()
    assert _rel(p8.grad[:, :4].cpu(), q.grad[:, :4]) < 2e-2
    del lat


def test_weights_preservation_and_gscale(cuda):
    from yat_b200.loss import fused_denoise_loss
    shape = (6, 32, 32, 32)
    pred, x0, noise, pf = _mk(shape, 1), _mk(shape, 2), _mk(shape, 3), _mk(shape, 4)
    w = torch.tensor([0.5, 1.0, 2.0, 0.0, 1.5, 0.25])
    total, l_noise, l_pres, g_ref = _oracle(pred, x0, noise, mode="velocity", pred_frozen=pf, rho=0.5, weights=w)
    p = pred.to(cuda).requires_grad_(True)
    loss, terms = fused_denoise_loss(p, x0.to(cuda), noise.to(cuda), pred_frozen=pf.to(cuda), weights=w.to(cuda),
                                     rho=0.5, return_terms=True)
    loss.backward()
    assert abs(loss.item() - total.item()) <= 1e-5 * abs(total.item())
    assert abs(terms[1].item() - l_noise.item()) <= 1e-5 * abs(l_noise.item())
    assert abs(terms[2].item() - l_pres.item()) <= 1e-5 * abs(l_pres.item())
    assert _rel(p.grad.cpu(), g_ref) < 4e-3
    # gradient accumulation: upstream gradient 1/4, both the trusted single-pass and the rescaled path
    for trust in (True, False):
        p2 = pred.to(cuda).requires_grad_(True)
        l2 = fused_denoise_loss(p2, x0.to(cuda), noise.to(cuda), gscale=0.25, trust_gscale=trust)
        (l2 / 4).backward()
        _, _, _, g1 = _oracle(pred, x0, noise, mode="velocity")
        assert _rel(p2.grad.cpu(), g1 * 0.25) < 4e-3


def test_regenerated_noise_matches_injected(cuda):
    """noise=None: eps is regenerated from (seed, offset) and must equal the noising kernel's eps."""
    from yat_b200.loss import fused_denoise_loss
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    shape = (4, 32, 32, 32)
    x0, pred = _mk(shape, 5).to(cuda), _mk(shape, 6).to(cuda)
    seed, offset = 0xABCDEF0123, 77
    nz = flowmatch_noise(x0, FlowMatchEulerDiscreteScheduler(), seed=seed, offset=offset, want_noise=True)
    p1 = pred.clone().requires_grad_(True)
    l1 = fused_denoise_loss(p1, x0, nz.noise)
    l1.backward()
    p2 = pred.clone().requires_grad_(True)
    l2 = fused_denoise_loss(p2, x0, None, seed=seed, offset=offset)
    l2.backward()
    assert torch.equal(l1, l2) and torch.equal(p1.grad, p2.grad)
    # and the regenerated stream is the oracle's stream (bf16 rounding flips allowed, rare)
    z = torch.from_numpy(philox.noise(seed, offset, 4, 32 * 32 * 32)).to(torch.bfloat16).view(shape)
    assert (z != nz.noise.cpu()).float().mean().item() < 2e-3


def test_bulk_properties(cuda):
    """> L2 slab: linearity in gscale, zero loss at pred == target, sum over shards == whole."""
    from yat_b200.loss import fused_denoise_loss
    B, C, H, W = 1024, 32, 32, 32
    g = torch.Generator(device=cuda).manual_seed(0)
    x0 = torch.randn((B, C, H, W), generator=g, device=cuda).to(torch.bfloat16)
    nz = torch.randn((B, C, H, W), generator=g, device=cuda).to(torch.bfloat16)
    tgt = nz - x0
    p = tgt.clone().requires_grad_(True)
    l0 = fused_denoise_loss(p, x0, nz)
    l0.backward()
    assert l0.item() == 0.0 and (p.grad == 0).all()
    pred = torch.randn((B, C, H, W), generator=g, device=cuda).to(torch.bfloat16)
    whole = fused_denoise_loss(pred, x0, nz)
    halves = [fused_denoise_loss(pred[i:i + B // 2].contiguous(), x0[i:i + B // 2].contiguous(),
                                 nz[i:i + B // 2].contiguous()) for i in (0, B // 2)]
    assert abs(whole.item() - 0.5 * (halves[0].item() + halves[1].item())) <= 1e-5 * whole.item()
    ref = torch.nn.functional.mse_loss(pred.float(), tgt.float())
    assert abs(whole.item() - ref.item()) <= 1e-5 * ref.item()


def test_argument_errors(cuda):
    from yat_b200.loss import fused_denoise_loss
    z = torch.zeros((2, 4, 3, 3), dtype=torch.bfloat16, device=cuda)   # chw = 36, not a multiple of 8
    with pytest.raises(ValueError):
        fused_denoise_loss(z, z, z)
    with pytest.raises(ValueError):
        fused_denoise_loss(z.float(), z, z)


@pytest.mark.parametrize("scheme", ["sigma_sqrt", "cosmap", "none"])
def test_sd3_weighting_scheme_through_the_hook_path(cuda, scheme):
    """weighting_scheme (config key) -> per-sample weights from the device sigmas -> yat_loss_fused(weights=), against the
    oracle restatement of diffusers' compute_loss_weighting_for_sd3 + the weighted per-sample MSE (oracle/losses.py)."""
    from oracle import losses as ol
    from yat_b200.loss import VELOCITY, fused_denoise_loss, sd3_loss_weights
    g = torch.Generator().manual_seed(7)
    B = 6
    x0, noise = [torch.randn((B, 16, 24, 24), generator=g).to(torch.bfloat16) for _ in range(2)]
    pred = torch.randn((B, 16, 24, 24), generator=g).to(torch.bfloat16)
    sig = torch.rand((B,), generator=g) * 0.9 + 0.05
    w_ref = ol.sd3_loss_weighting(scheme, sig)
    pr = pred.clone().float().requires_grad_(True)
    loss_ref = ol.denoise_loss(pr, x0, noise, mode="velocity", weights=w_ref)
    loss_ref = loss_ref[0] if isinstance(loss_ref, tuple) else loss_ref
    loss_ref.backward()
    w = sd3_loss_weights(scheme, sig.to(cuda))
    assert (w is None) == (scheme == "none")
    pc = pred.to(cuda).requires_grad_(True)
    loss = fused_denoise_loss(pc, x0.to(cuda), noise.to(cuda), weights=w, target_mode=VELOCITY, lowp=False)
    loss.backward()
    assert abs(loss.item() - loss_ref.item()) <= 1e-5 * abs(loss_ref.item())
    num = (pc.grad.float().cpu() - pr.grad).norm() / pr.grad.norm()
    assert num < 2e-2
