"""yat_gemm_bf16_tn (tcgen05, cta_group::2, TMA, TMEM double-buffered accumulators) against torch: the fp64 product of
the same bf16 operands (truth) and the library bf16 GEMM (what F.linear runs, utils/patch_sana_attention_layers.py:84-113
projections).  Tolerance: bf16 output, fp32 accumulation -> one rounding of the exact result up to summation order:
<= 2e-2 relative by spec, asserted at the bf16 rounding level (4e-3) and not further from fp64 than cuBLAS."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm().clamp_min(1e-300)).item()


# (M, N, K, tile_n): the SANA-1.6B projection shapes at reduced M, plus ragged M / N / K tails and every tile width
SHAPES = [(512, 6720, 2240, 0), (512, 6720, 2240, 256), (1024, 2240, 2240, 0), (300, 11200, 2240, 0),
          (256, 2240, 5600, 0), (130, 96, 72, 128), (257, 500, 136, 256), (1, 32, 8, 128), (777, 224, 64, 224),
          (2048, 448, 4096, 224)]


@pytest.mark.parametrize("M,N,K,tile_n", SHAPES)
@pytest.mark.parametrize("with_bias", [False, True])
def test_gemm_matches_fp64_and_library(cuda, M, N, K, tile_n, with_bias):
    from yat_b200.gemm import gemm_bf16_tn
    g = torch.Generator(device=cuda).manual_seed(M * 7 + N + K)
    a = torch.randn((M, K), device=cuda, generator=g).to(torch.bfloat16)
    w = (torch.randn((N, K), device=cuda, generator=g) / K ** 0.5).to(torch.bfloat16)
    b = torch.randn((N,), device=cuda, generator=g).to(torch.bfloat16) if with_bias else None
    c = gemm_bf16_tn(a, w, b, tile_n=tile_n)
    assert c.shape == (M, N) and c.dtype == torch.bfloat16
    ref = a.double() @ w.double().t()
    if with_bias:
        ref = ref + b.double()
    lib = torch.nn.functional.linear(a, w, b)
    e_own, e_lib = _rel(c, ref), _rel(lib, ref)
    assert e_own < 4e-3, (e_own, e_lib)
    assert e_own <= 1.05 * e_lib + 1e-5, (e_own, e_lib)
    # elementwise: within one bf16 ulp of the rounded exact result (fp32 accumulation order is the only freedom)
    exact = ref.to(torch.bfloat16).float()
    assert ((c.float() - exact).abs() <= 2.0 ** -7 * exact.abs() + 1e-6 + 2e-3 * ref.abs().mean()).all()
    # run-to-run identical (fixed accumulation order: K blocks in sequence, one accumulator per tile)
    assert torch.equal(c, gemm_bf16_tn(a, w, b, tile_n=tile_n))


def test_gemm_strided_views_and_batched_leading_dims(cuda):
    """a as a column slice of a wider matrix (row pitch > K), a leading batch dimension, out= into a pitched buffer."""
    from yat_b200.gemm import gemm_bf16_tn
    g = torch.Generator(device=cuda).manual_seed(5)
    wide = torch.randn((4, 96, 3 * 256), device=cuda, generator=g).to(torch.bfloat16)
    a = wide[..., 256:512]                                         # pitch 768, offset 256 elements
    w = (torch.randn((320, 256), device=cuda, generator=g) / 16).to(torch.bfloat16)
    c = gemm_bf16_tn(a, w)
    assert c.shape == (4, 96, 320)
    assert _rel(c, a.double() @ w.double().t()) < 4e-3
    out = torch.full((4 * 96, 3 * 320), 7.0, device=cuda, dtype=torch.bfloat16)
    c2 = gemm_bf16_tn(a, w, out=out[:, 320:640])
    assert torch.equal(out[:, 320:640].reshape(4, 96, 320), c) and torch.equal(c2, c)
    assert (out[:, :320] == 7).all() and (out[:, 640:] == 7).all()  # nothing written outside the tile columns


def test_gemm_full_size_qkv_projection(cuda):
    """The SANA-1.6B fused q|k|v projection [16384, 2240] x [6720, 2240]^T at full size: size-independent properties
    (linearity in a, a column block equals the GEMM on that block of rows of w) + agreement with the library."""
    from yat_b200.gemm import gemm_bf16_tn
    g = torch.Generator(device=cuda).manual_seed(9)
    a = torch.randn((16384, 2240), device=cuda, generator=g).to(torch.bfloat16)
    w = (torch.randn((6720, 2240), device=cuda, generator=g) / 47.0).to(torch.bfloat16)
    c = gemm_bf16_tn(a, w)
    lib = torch.nn.functional.linear(a, w)
    assert _rel(c, lib.float()) < 3e-3
    assert torch.equal(c[:, 2240:4480], gemm_bf16_tn(a, w[2240:4480]))       # columns depend only on their own w rows
    c2 = gemm_bf16_tn((a.float() * 2).to(torch.bfloat16), w)                  # exact power-of-two scaling
    assert torch.equal(c2.float(), 2 * c.float())
    rows = torch.randint(0, 16384, (64,), device=cuda, generator=g)
    ref = a[rows].double() @ w.double().t()
    assert _rel(c[rows], ref) < 3e-3


def test_gemm_argument_errors(cuda):
    from yat_b200.gemm import gemm_bf16_tn
    a = torch.zeros((8, 16), device=cuda, dtype=torch.bfloat16)
    with pytest.raises(ValueError):
        gemm_bf16_tn(a, torch.zeros((8, 24), device=cuda, dtype=torch.bfloat16))
    with pytest.raises(ValueError):
        gemm_bf16_tn(a.float(), torch.zeros((8, 16), device=cuda, dtype=torch.bfloat16))
    with pytest.raises(ValueError):
        gemm_bf16_tn(a.cpu(), torch.zeros((8, 16), dtype=torch.bfloat16))
    with pytest.raises(ValueError):
        gemm_bf16_tn(a, torch.zeros((8, 16), device=cuda, dtype=torch.bfloat16), tile_n=192)
