"""GPU: fused noising kernel vs the CPU oracle, through the C-ABI (yat_noise_flowmatch / _ddpm)."""
import hashlib

import numpy as np
import pytest
import torch

from oracle import noising, philox, sampling, schedule

pytestmark = pytest.mark.gpu


def sha(t):
    return hashlib.sha256(t.contiguous().cpu().view(torch.uint8).numpy().tobytes()).hexdigest()


def _draws(shape, dtype, seed=None):
    g = torch.Generator() if seed is None else torch.Generator().manual_seed(seed)
    return sampling.sana_step_draws(shape, shape[0], dtype, g)


def test_k2_inject_mode_is_bit_exact_and_matches_golden(cuda, kat):
    """SANA-1.6B per-GPU batch [16,32,32,32] bf16 with the reference's own fresh-generator draws."""
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    noise, u, idx = _draws((16, 32, 32, 32), torch.bfloat16)
    lat = torch.randn((16, 32, 32, 32), generator=torch.Generator().manual_seed(1234)).to(torch.bfloat16)
    tt, ss = schedule.flow_match_tables(1000, 1.0)
    x_ref, tg_ref, t_ref, sig_ref = noising.flow_noise_sana(lat, noise, idx, tt, ss)
    out = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(), noise=noise.to(cuda), u=u.to(cuda))
    assert out.idx.cpu().tolist() == idx.tolist() == kat["K2"]["idx"]          # bit-exact indices
    assert torch.equal(out.timesteps.cpu(), t_ref)                                # bit-exact timesteps
    assert torch.equal(out.sigmas.cpu(), sig_ref.flatten().float())               # sigma after the bf16 cast
    assert torch.equal(out.x_t.cpu(), x_ref) and torch.equal(out.target.cpu(), tg_ref)
    assert sha(out.x_t) == kat["K2"]["x_t_sha256"] and sha(out.target) == kat["K2"]["target_sha256"]


@pytest.mark.parametrize("shape,shift", [((2, 32, 16, 16), 1.0), ((8, 16, 128, 128), 3.0), ((16, 4, 128, 128), 1.0),
                                         ((3, 4, 88, 176), 3.0), ((5, 3, 7, 9), 1.0)])
def test_flow_inject_bf16_shapes(cuda, shape, shift):
    """K1/K4/K3 latent shapes, a non-square bucket, and a ragged shape (chw % 8 != 0 -> generic kernel)."""
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    noise, u, idx = _draws(shape, torch.bfloat16, seed=42)
    lat = torch.randn(shape, generator=torch.Generator().manual_seed(7)).to(torch.bfloat16)
    tt, ss = schedule.flow_match_tables(1000, shift)
    x_ref, tg_ref, t_ref, _ = noising.flow_noise_sd3(lat, noise, idx, tt, ss)
    out = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(shift=shift), noise=noise.to(cuda), u=u.to(cuda))
    assert out.idx.cpu().tolist() == idx.tolist()
    assert torch.equal(out.timesteps.cpu(), t_ref)
    assert torch.equal(out.x_t.cpu(), x_ref) and torch.equal(out.target.cpu(), tg_ref)


def test_k1_fp32_inject(cuda, kat):
    """Config K1 arithmetic (fp32 latents): no FMA contraction, bit-exact against torch CPU."""
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    noise, u, idx = _draws((2, 32, 16, 16), torch.float32)
    lat = torch.randn((2, 32, 16, 16), generator=torch.Generator().manual_seed(1))
    tt, ss = schedule.flow_match_tables(1000, 1.0)
    x_ref, tg_ref, t_ref, _ = noising.flow_noise_sana(lat, noise, idx, tt, ss)
    out = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(), noise=noise.to(cuda), u=u.to(cuda))
    assert out.idx.cpu().tolist() == idx.tolist()
    assert torch.equal(out.x_t.cpu(), x_ref) and torch.equal(out.target.cpu(), tg_ref)


def test_ddpm_inject_bit_exact(cuda, kat):
    """PixArt-Sigma: DDPM add_noise, eps target, int64 timesteps = T-1-idx."""
    from yat_b200.noise import ddpm_noise
    from yat_b200.schedulers import DDPMScheduler
    for shape in ((16, 32, 32, 32), (4, 4, 128, 128), (2, 4, 11, 13)):
        noise, u, idx = _draws(shape, torch.bfloat16)
        lat = torch.randn(shape, generator=torch.Generator().manual_seed(1234)).to(torch.bfloat16)
        dt, acp = schedule.ddpm_tables()
        x_ref, _, t_ref, _ = noising.ddpm_noise(lat, noise, idx, dt, acp)
        sch = DDPMScheduler()
        out = ddpm_noise(lat.to(cuda), sch, noise=noise.to(cuda), u=u.to(cuda))
        assert out.timesteps.cpu().tolist() == t_ref.tolist()
        assert torch.equal(out.x_t.cpu(), x_ref)
        if shape == (16, 32, 32, 32):
            assert sha(out.x_t) == kat["K2"]["ddpm_x_t_sha256"]
            assert out.timesteps[:4].cpu().tolist() == [232, 154, 320, 360]
        # diffusers-style entry point: add_noise(samples, noise, timesteps)
        assert torch.equal(sch.add_noise(lat.to(cuda), noise.to(cuda), t_ref.to(cuda)).cpu(), x_ref)


def test_index_edge_cases(cuda):
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    u = torch.tensor([1.0, 0.99999994, 0.0, 0.5, 0.001, 0.0009999], dtype=torch.float32)
    lat = torch.zeros((6, 8), dtype=torch.bfloat16)
    out = flowmatch_noise(lat.to(cuda), FlowMatchEulerDiscreteScheduler(), noise=lat.to(cuda), u=u.to(cuda))
    assert out.idx.cpu().tolist() == sampling.timestep_indices(u, 1000).tolist() == [999, 999, 0, 500, 1, 0]


def test_generate_mode_matches_the_philox_oracle(cuda):
    from yat_b200 import _lib
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    seed, offset, B, chw = 0x1234567890ABCDEF, (5 << 32) + 17, 3, 4096
    # integer stream: bit-exact
    raw = torch.empty((64, 4), dtype=torch.int32, device=cuda)
    for stream in (0, 1):
        rc = _lib.lib().yat_philox_raw(seed, offset, 2, stream, 64, _lib.ptr(raw), _lib.current_stream())
        _lib.check(rc, "yat_philox_raw")
        assert np.array_equal(raw.cpu().numpy().view(np.uint32), philox.raw_stream(seed, offset, 2, 64, stream))
    lat = torch.randn((B, chw), generator=torch.Generator().manual_seed(3)).to(torch.bfloat16)
    sch = FlowMatchEulerDiscreteScheduler()
    out = flowmatch_noise(lat.to(cuda), sch, seed=seed, offset=offset, want_noise=True)
    # normals: ~1e-6 absolute from the float64 oracle -> after bf16 rounding at most rare 1-ulp flips
    z = torch.from_numpy(philox.noise(seed, offset, B, chw))
    got = out.noise.cpu().float()
    ref = z.to(torch.bfloat16).float()
    mism = (got != ref)
    assert mism.float().mean().item() < 2e-3
    # a flip is one bf16 ulp, or (values near zero, where sin/cos.approx are absolute-accurate) < 4e-6
    assert (((got - ref).abs() <= 2.0 ** -7 * ref.abs() * 1.01) | ((got - ref).abs() < 4e-6))[mism].all()
    assert abs(got.mean().item()) < 0.02 and abs(got.std().item() - 1.0) < 0.02
    # timestep indices: equal to the oracle unless u*T sits on an integer boundary
    zt = philox.timestep_normals(seed, offset, B)
    u = 1.0 / (1.0 + np.exp(-zt))
    idx_ref = np.minimum((u * 1000).astype(np.int64), 999)
    near = np.abs(u * 1000 - np.round(u * 1000)) < 1e-3
    idx = out.idx.cpu().numpy()
    assert np.all((idx == idx_ref) | near)
    tt, ss = schedule.flow_match_tables(1000, 1.0)
    assert torch.equal(out.timesteps.cpu(), tt[out.idx.cpu()])
    # given the generated eps and idx, x_t/target follow the reference arithmetic exactly
    x_ref, tg_ref, _, _ = noising.flow_noise_sana(lat, out.noise.cpu(), out.idx.cpu(), tt, ss)
    assert torch.equal(out.x_t.cpu(), x_ref) and torch.equal(out.target.cpu(), tg_ref)
    # deterministic, and offset/seed/rank-sensitive
    again = flowmatch_noise(lat.to(cuda), sch, seed=seed, offset=offset, want_noise=True)
    assert torch.equal(again.noise, out.noise) and torch.equal(again.x_t, out.x_t)
    other = flowmatch_noise(lat.to(cuda), sch, seed=seed, offset=offset + 1, want_noise=True)
    assert not torch.equal(other.noise, out.noise)
    # ragged chw goes through the generic kernel and must draw the same stream
    lat2 = lat[:, :4093].contiguous()
    o2 = flowmatch_noise(lat2.to(cuda), sch, seed=seed, offset=offset, want_noise=True)
    assert torch.equal(o2.noise.cpu(), out.noise.cpu()[:, :4093])


def test_full_size_properties(cuda):
    """Bulk slab (512 MiB of latents, > L2): size-independent properties instead of a CPU compare."""
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    B, chw = 2048, 32 * 32 * 32
    g = torch.Generator(device=cuda).manual_seed(0)
    lat = torch.randn((B, chw), generator=g, device=cuda, dtype=torch.float32).to(torch.bfloat16)
    sch = FlowMatchEulerDiscreteScheduler()
    # sigma == 1 (idx 0): x_t == eps exactly, target == eps - x0
    u0 = torch.zeros(B, device=cuda)
    o = flowmatch_noise(lat, sch, u=u0, seed=9, offset=1, want_noise=True)
    assert torch.equal(o.x_t, o.noise) and (o.sigmas == 1).all()
    assert torch.equal(o.target, (o.noise - lat))
    # the whole batch equals the per-sample op (samples are independent: the sharding property)
    o_full = flowmatch_noise(lat, sch, seed=9, offset=2, want_noise=True)
    half = flowmatch_noise(lat[: B // 2].contiguous(), sch, seed=9, offset=2, want_noise=True)
    assert torch.equal(o_full.x_t[: B // 2], half.x_t) and torch.equal(o_full.idx[: B // 2], half.idx)
    # x0 == 0 -> x_t == bf16(sigma * eps), target == eps
    z = torch.zeros_like(lat)
    oz = flowmatch_noise(z, sch, seed=9, offset=2, want_noise=True)
    assert torch.equal(oz.target, oz.noise)
    assert torch.equal(oz.x_t, (oz.sigmas.to(torch.bfloat16)[:, None] * oz.noise))
    # idx histogram is logit-normal(0,1): median near T/2
    med = oz.idx.float().median().item()
    assert 440 < med < 560


def test_argument_errors(cuda):
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    sch = FlowMatchEulerDiscreteScheduler()
    lat = torch.zeros((2, 16), dtype=torch.bfloat16, device=cuda)
    with pytest.raises(ValueError):
        flowmatch_noise(lat, sch, noise=torch.zeros((2, 8), dtype=torch.bfloat16, device=cuda))
    with pytest.raises(ValueError):
        flowmatch_noise(lat.to(torch.float16), sch)
    with pytest.raises(ValueError):
        flowmatch_noise(lat, sch, u=torch.zeros(3, device=cuda))


def test_generate_mode_noise_is_standard_normal_on_the_bf16_grid(cuda):
    """The stream every real training step uses (generate mode, train_sana.py:183 draws eps ~ N(0,1) in bf16):
    2^27 bf16 samples of the counter RNG against the bf16-ROUNDED standard normal (oracle/normal_grid.py: exact bin
    masses from the float64 normal CDF).  Moments within 3.5 sampling standard errors, discrete
    Kolmogorov-Smirnov distance below the alpha = 0.001 critical value, tail masses beyond 3, 4 and 4.5 sigma,
    lag-1 and within-pair correlations (the two normals of a pair come from ONE Philox word), per-sample streams
    uncorrelated.  Sampling errors: mean 1/sqrt(n), variance sqrt(2/n), skewness sqrt(6/n), kurtosis sqrt(24/n)."""
    from oracle import normal_grid
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    B, chw = 4096, 32 * 32 * 32
    n = B * chw
    assert n == 1 << 27
    lat = torch.zeros((B, chw), dtype=torch.bfloat16, device=cuda)
    out = flowmatch_noise(lat, FlowMatchEulerDiscreteScheduler(), seed=20261104, offset=3, want_target=False,
                          want_noise=True)
    z = out.noise
    del lat
    # ---- histogram over the 65536 bf16 patterns (exact empirical distribution)
    counts = torch.zeros(65536, dtype=torch.int64, device=cuda)
    for part in z.view(torch.int16).reshape(-1).chunk(16):
        counts += torch.bincount(part.to(torch.int32) & 0xFFFF, minlength=65536)
    counts = counts.cpu().numpy()
    assert counts.sum() == n
    vals, bits, mass, cdf = normal_grid.bf16_normal_bins()
    assert counts[~np.isin(np.arange(65536), np.concatenate([bits, [0x8000]]))].sum() == 0      # no inf / nan
    d, _, at = normal_grid.discrete_ks(counts)
    ks_crit = 1.95 / np.sqrt(n)                                # Kolmogorov, alpha = 0.001
    assert d < ks_crit, (d, ks_crit, at)
    # ---- moments of the rounded variable vs their exact expectations
    c = counts.astype(np.float64).copy()
    c[0] += c[0x8000]
    p = c[bits] / n
    m1 = float((p * vals).sum())
    mu2 = float((p * (vals - m1) ** 2).sum())
    skew = float((p * (vals - m1) ** 3).sum()) / mu2 ** 1.5
    kurt = float((p * (vals - m1) ** 4).sum()) / mu2 ** 2 - 3.0
    e1, e2, e3, e4 = normal_grid.rounded_normal_moments()
    k = 3.5
    assert abs(m1 - e1) < k / np.sqrt(n), m1
    assert abs(mu2 - e2) < k * np.sqrt(2.0 / n), mu2
    assert abs(skew - e3) < k * np.sqrt(6.0 / n), skew
    assert abs(kurt - e4) < k * np.sqrt(24.0 / n), kurt
    # ---- tail masses (two-sided) vs the rounded normal: binomial standard error
    for t in (3.0, 4.0, 4.5):
        exp = float(mass[np.abs(vals) >= t].sum())
        got = float(p[np.abs(vals) >= t].sum())
        se = np.sqrt(exp * (1 - exp) / n)
        assert abs(got - exp) < 4 * se, (t, got, exp, se)
    # the one-word Box-Muller stops at sqrt(40 ln 2) = 5.27 (torch's fp32 CPU randn at 5.77): the mass beyond is
    # 1.4e-7, i.e. 18 of 2^27 samples
    assert np.abs(vals[p > 0]).max() <= 5.3
    # ---- correlations: lag 1 inside a sample (element 2j, 2j+1 share a Philox word; 2j+1, 2j+2 do not),
    # across samples at the same element, squares (a shared-radius signature) - all ~ N(0, 1/sqrt(m))
    zf = z[:512].float()                                       # 2^24 elements
    m = zf.numel()
    def corr(a, b):
        a, b = a.reshape(-1).double(), b.reshape(-1).double()
        return float(((a - a.mean()) * (b - b.mean())).mean() / (a.std() * b.std()))
    pair = corr(zf[:, 0::2], zf[:, 1::2])
    lag = corr(zf[:, 1:-1:2], zf[:, 2::2])
    cross = corr(zf[:256], zf[256:])
    sq_lag = corr(zf[:, 1:-1:2] ** 2, zf[:, 2::2] ** 2)
    lim = 4.5 / np.sqrt(m / 2)
    assert abs(pair) < lim and abs(lag) < lim and abs(cross) < lim and abs(sq_lag) < lim, (pair, lag, cross, sq_lag)
    # the two normals of one word share the radius: their squares ARE correlated in any Box-Muller pair only if
    # the angle were not uniform; with a uniform angle corr(z0^2, z1^2) = 0 exactly (chi-square_2 radius)
    sq_pair = corr(zf[:, 0::2] ** 2, zf[:, 1::2] ** 2)
    assert abs(sq_pair) < lim, sq_pair
