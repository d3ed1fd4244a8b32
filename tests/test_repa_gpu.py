"""GPU: REPA cosine-alignment kernel (yat_repa_cos) vs the CPU oracle (forward + gradient)."""
import pytest
import torch

from oracle import repa as orepa

pytestmark = pytest.mark.gpu


def _case(B, gh, gw, E, y_dtype, seed=0):
    g = torch.Generator().manual_seed(seed)
    z = torch.randn((B, gh * gw, E), generator=g).to(torch.bfloat16)
    y = torch.nn.functional.normalize(torch.randn((B, 256, E), generator=g), dim=-1).to(y_dtype)
    return z, y


def _oracle(z, y, grid):
    zz = z.float().requires_grad_(True)
    loss = orepa.repa_cosine_loss(zz, y, grid)
    loss.backward()
    return loss.detach(), zz.grad


@pytest.mark.parametrize("B,gh,gw,E,y_dtype", [
    (16, 32, 32, 768, torch.float32),    # K5: SANA 1024px square bucket, DINOv2 fp32 teacher
    (4, 32, 32, 768, torch.bfloat16),
    (3, 22, 44, 768, torch.float32),     # non-square bucket: overlapping adaptive windows
    (2, 16, 64, 768, torch.bfloat16),
    (2, 64, 64, 768, torch.float32),     # PixArt 64x64 token grid -> 4x4 windows
    (2, 16, 16, 256, torch.float32),     # identity pooling, small E
    (2, 20, 24, 1024, torch.float32),
])
def test_forward_and_gradient(cuda, B, gh, gw, E, y_dtype):
    from yat_b200.repa import repa_cosine_loss
    z, y = _case(B, gh, gw, E, y_dtype)
    l_ref, g_ref = _oracle(z, y, (gh, gw))
    zc = z.to(cuda).requires_grad_(True)
    loss = repa_cosine_loss(zc, y.to(cuda), (gh, gw))
    loss.backward()
    # fp32 loss: 1e-5 relative, with an absolute floor because the mean cosine of random vectors is ~0
    assert abs(loss.item() - l_ref.item()) <= 1e-5 * abs(l_ref.item()) + 5e-7
    rel = ((zc.grad.cpu().float() - g_ref).norm() / g_ref.norm()).item()
    assert rel < 2e-2, rel      # dz is stored in bf16
    assert rel < 4e-3           # in practice it is one bf16 rounding


def test_aligned_tokens_give_minus_one(cuda):
    from yat_b200.repa import repa_cosine_loss
    g = torch.Generator().manual_seed(1)
    y = torch.randn((2, 256, 768), generator=g)
    # student = teacher upsampled 2x2 (so the pooled student equals the teacher up to bf16 rounding)
    z = y.view(2, 16, 16, 768).repeat_interleave(2, 1).repeat_interleave(2, 2).reshape(2, 1024, 768)
    loss = repa_cosine_loss(z.to(torch.bfloat16).to(cuda), y.to(cuda), (32, 32))
    assert abs(loss.item() + 1.0) < 1e-4


def test_upstream_gradient_scales(cuda):
    from yat_b200.repa import repa_cosine_loss
    z, y = _case(2, 32, 32, 768, torch.float32, seed=3)
    a = z.to(cuda).requires_grad_(True)
    repa_cosine_loss(a, y.to(cuda), (32, 32)).backward()
    b = z.to(cuda).requires_grad_(True)
    (0.05 * repa_cosine_loss(b, y.to(cuda), (32, 32))).backward()
    rel = ((b.grad.float() - 0.05 * a.grad.float()).norm() / (0.05 * a.grad.float()).norm()).item()
    assert rel < 8e-3


def test_repa_model_wrapper_stashes_and_trains(cuda):
    """RepaModel/RepaModule surface (common/repa.py:59-135) on a toy base model."""
    from yat_b200.repa import RepaConfig, RepaModel

    class Base(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.blocks = torch.nn.ModuleList([torch.nn.Linear(64, 64) for _ in range(2)])

        def forward(self, x):
            for b in self.blocks:
                x = b(x)
            return x

    base = Base().to(cuda).to(torch.bfloat16)
    m = RepaModel(base, RepaConfig(target_modules=["blocks.0"], target_shape=768, hidden_shape=[64]))
    x = torch.randn((2, 1024, 64), device=cuda, dtype=torch.bfloat16)
    y = torch.nn.functional.normalize(torch.randn((2, 256, 768), device=cuda), dim=-1)
    out = m(x)
    assert out.shape == x.shape and m.model.get_submodule("blocks.0").zs.shape == (2, 1024, 768)
    loss = m.calculate_loss(y, (32, 32))
    loss.backward()
    head = m.model.get_submodule("blocks.0").proj_mlp
    assert head.fc2.weight.grad is not None and head.fc2.weight.grad.abs().sum() > 0
    assert m.model.get_submodule("blocks.0").zs is None


def test_argument_errors(cuda):
    from yat_b200.repa import repa_cosine_loss
    z = torch.zeros((2, 1024, 768), dtype=torch.bfloat16, device=cuda)
    y = torch.zeros((2, 256, 768), device=cuda)
    with pytest.raises(ValueError):
        repa_cosine_loss(z, y, (16, 16))
    with pytest.raises(ValueError):
        repa_cosine_loss(z.float(), y, (32, 32))
