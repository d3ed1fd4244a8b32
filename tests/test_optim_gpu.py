"""Clip + AdamW kernels (yat_grad_clip_coef / yat_adamw_step) vs the oracle restatement of
common/trainer.py:347-348 (clip_grad_norm_ + torch.optim.AdamW.step)."""
import pytest
import torch

from oracle import optimizer as oopt

pytestmark = pytest.mark.gpu

SIZES = [(7,), (2240,), (6, 2240), (65536,), (65537,), (300, 1001), (2240, 2240), (3,), (131072 + 8,)]


def _make(dtype, seed, gscale):
    g = torch.Generator().manual_seed(seed)
    ps = [torch.randn(s, generator=g).to(dtype) for s in SIZES]
    gs = [(torch.randn(s, generator=g) * gscale).to(dtype) for s in SIZES]
    return ps, gs


def _ulp_close(got, ref, dtype, max_flip_frac, scale=None):
    """Equal up to one rounding flip of the stored dtype.  ``scale`` = magnitude of the operands the value
    was formed from: where the update cancels the parameter (p ~ lr) the fp32 op chain is exact to
    ~1e-6 of the operands, not of the (tiny) result."""
    got, ref = got.float().cpu(), ref.float()
    mism = got != ref
    # bf16: one ulp; fp32: the fp32 op chain of the kernel vs one rounding of the float64 result
    ulp = (2.0 ** -7 if dtype == torch.bfloat16 else 2.0 ** -20) * ref.abs().clamp_min(1e-30)
    if scale is not None:
        ulp = torch.maximum(ulp, 2e-6 * scale.float())
    assert ((got - ref).abs() <= ulp * 1.01)[mism].all(), (got - ref).abs().max()
    assert mism.sum().item() <= max(1, max_flip_frac * mism.numel()), mism.float().mean().item()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("gscale", [1e-4, 1e-2])   # total norm below / above max_norm = 1
def test_clip_adamw_matches_oracle(cuda, dtype, gscale):
    from yat_b200.optim import FusedClipAdamW
    lr, wd = 1e-2, 0.05
    ps, gs = _make(dtype, 5, gscale)
    dev_ps = [torch.nn.Parameter(p.to(cuda)) for p in ps]
    opt = FusedClipAdamW(dev_ps, lr=lr, weight_decay=wd)
    ref_ps, ref_state = [p.clone() for p in ps], None
    for step in range(1, 4):
        # fresh gradients every step (scaled copies of one draw make m = 0.9 m + 0.1 g land on exact bf16 ties)
        gen = torch.Generator().manual_seed(100 + step)
        grads = [(torch.randn(g.shape, generator=gen) * gscale * (1.0 + 0.25 * step)).to(dtype) for g in gs]
        for p, g in zip(dev_ps, grads):
            p.grad = g.to(cuda)
        # the oracle starts every step from the device's own stored state, so one-ulp rounding flips of
        # an earlier step do not compound into the comparison
        cur = [p.detach().cpu().clone() for p in dev_ps]
        m = [opt.state[p]["exp_avg"].cpu().clone() if len(opt.state[p]) else torch.zeros_like(q) for p, q in zip(dev_ps, ps)]
        v = [opt.state[p]["exp_avg_sq"].cpu().clone() if len(opt.state[p]) else torch.zeros_like(q) for p, q in zip(dev_ps, ps)]
        opt.step(max_norm=1.0)
        p_in, m_in, v_in = [t.clone() for t in cur], [t.clone() for t in m], [t.clone() for t in v]
        cur, m, v, norm, coef = oopt.exact_step(cur, grads, m, v, step, lr, wd, max_norm=1.0)
        ref_ps, ref_norm, ref_state = oopt.reference_step(ps, grads, ref_state, lr, wd, max_norm=1.0)
        assert abs(opt.last_norm.item() - norm) <= 1e-5 * norm
        assert (coef < 1.0) == (gscale == 1e-2)
        # same arithmetic evaluated in float64: identical up to rare one-ulp rounding flips
        # bf16 inputs have 256 distinct mantissas: one of them landing 0.001 g^2 on a bf16 midpoint already flips 0.6 %
        flips = 3e-2 if dtype == torch.bfloat16 else 1.0   # fp32 storage: only the error bound is meaningful
        for p, q, pi in zip(dev_ps, cur, p_in):
            _ulp_close(p.detach(), q, dtype, flips, scale=pi.abs() + lr)
        for p, a, b, mi, vi, g in zip(dev_ps, m, v, m_in, v_in, grads):
            _ulp_close(opt.state[p]["exp_avg"], a, dtype, flips, scale=mi.abs() + g.abs())      # 0.9 m + 0.1 g can cancel
            _ulp_close(opt.state[p]["exp_avg_sq"], b, dtype, flips, scale=vi.abs() + g.float() ** 2)
        # the reference's own torch calls (per-op rounding in the storage dtype): bf16 tolerance of the spec
        for p0, p, r in zip(ps, dev_ps, ref_ps):
            du, dr = p.detach().float().cpu() - p0.float(), r.float() - p0.float()
            tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
            # updates below the storage resolution of p are rounded away on both sides: compare what moved
            assert (du - dr).norm() <= tol * dr.norm() + (2.0 ** -8 if dtype == torch.bfloat16 else 1e-7) * p0.float().norm()
    assert int(opt.state[dev_ps[0]]["step"]) == 3


def test_plain_step_and_misaligned_views(cuda):
    """No clip (coef pointer NULL) and parameters / gradients that are odd-offset views of flat buffers
    (DDP bucket views): the scalar path must give the same numbers."""
    from yat_b200.optim import FusedClipAdamW
    g = torch.Generator().manual_seed(9)
    n = 70001
    flat_p = torch.randn(n + 3, generator=g).to(torch.bfloat16)
    flat_g = (torch.randn(n + 3, generator=g) * 0.1).to(torch.bfloat16)
    p_view = torch.nn.Parameter(flat_p.to(cuda)[3:])           # 6-byte offset: not 16-byte aligned
    p_view.grad = flat_g.to(cuda)[1:n + 1]
    p_al = torch.nn.Parameter(flat_p[3:].clone().to(cuda))
    p_al.grad = flat_g[1:n + 1].clone().to(cuda)
    for p in (p_view, p_al):
        FusedClipAdamW([p], lr=1e-2, weight_decay=0.0).step()
    assert torch.equal(p_view.detach(), p_al.detach())
    q, _, _, _, _ = oopt.exact_step([flat_p[3:]], [flat_g[1:n + 1]], [torch.zeros(n, dtype=torch.bfloat16)],
                                    [torch.zeros(n, dtype=torch.bfloat16)], 1, 1e-2, 0.0, max_norm=None)
    _ulp_close(p_al.detach(), q[0], torch.bfloat16, 3e-2, scale=flat_p[3:].abs() + 1e-2)


def test_deterministic_and_state_dict(cuda):
    from yat_b200.optim import FusedClipAdamW
    ps, gs = _make(torch.bfloat16, 11, 1e-2)
    outs = []
    for _ in range(2):
        dev = [torch.nn.Parameter(p.to(cuda)) for p in ps]
        opt = FusedClipAdamW(dev, lr=1e-3)
        for p, g in zip(dev, gs):
            p.grad = g.to(cuda)
        opt.clip_and_step(1.0)
        outs.append(([p.detach().clone() for p in dev], opt.last_norm.clone()))
    assert all(torch.equal(a, b) for a, b in zip(outs[0][0], outs[1][0])) and torch.equal(outs[0][1], outs[1][1])
    sd = opt.state_dict()
    ref = torch.optim.AdamW([torch.nn.Parameter(p.clone()) for p in ps], lr=1e-3)
    assert set(sd["param_groups"][0]) >= {"lr", "betas", "eps", "weight_decay", "params"}
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"}
    ref.load_state_dict(sd)   # layout-compatible with torch.optim.AdamW


# ---------------------------------------------------------------------------------------------------------
# EMA (common/trainer.py:264-268,350-351): yat_ema_step / yat_adamw_ema_step vs oracle/ema.py (torch ops on CPU)
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_ema_standalone_bit_exact_vs_oracle(cuda, dtype):
    """EMAModel.step alone: the shadow tensors equal the reference's three torch ops bit for bit (the kernel keeps
    the per-op rounding to the storage dtype), over the decay warm-up (1 - decay = 1, 9/11, 9/12, ...) and one
    frozen tensor (copied)."""
    from oracle.ema import EMAOracle, get_decay
    from yat_b200.ema import EMAModel
    ps, _ = _make(dtype, 11, 1.0)
    cpu = [torch.nn.Parameter(p.clone()) for p in ps]
    dev = [torch.nn.Parameter(p.to(cuda)) for p in ps]
    cpu[3].requires_grad_(False), dev[3].requires_grad_(False)
    ref, ema = EMAOracle(cpu, decay=0.999), EMAModel(dev, decay=0.999)
    gen = torch.Generator().manual_seed(3)
    for step in range(1, 7):
        for a, b in zip(cpu, dev):                       # the weights move between EMA steps
            d = (torch.randn(a.shape, generator=gen) * 0.05).to(dtype)
            a.data.add_(d)
            b.data.add_(d.to(cuda))
        omd = ref.step(cpu)
        ema.step(dev)
        assert ema.optimization_step == step and abs((1 - ema.cur_decay_value) - omd) < 1e-15
        assert ema.cur_decay_value == get_decay(step)
        for s_ref, s in zip(ref.shadow_params, ema.shadow_params):
            assert torch.equal(s.cpu(), s_ref), (step, s.shape)
    assert ema.cur_decay_value == pytest.approx(6 / 15)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_ema_fused_into_adamw_equals_separate_steps(cuda, dtype):
    """optimizer.clip_and_step(ema=...) == optimizer.clip_and_step(); ema.step(parameters): weights, moments and
    shadows bit-identical (the fused kernel feeds the EMA the value it just stored), and the shadow matches the
    oracle's torch ops applied to the device's own weights."""
    from oracle.ema import EMAOracle
    from yat_b200.ema import EMAModel
    from yat_b200.optim import FusedClipAdamW
    ps, gs = _make(dtype, 21, 1e-2)
    a = [torch.nn.Parameter(p.to(cuda)) for p in ps]
    b = [torch.nn.Parameter(p.to(cuda)) for p in ps]
    oa, ob = FusedClipAdamW(a, lr=1e-2, weight_decay=0.05), FusedClipAdamW(b, lr=1e-2, weight_decay=0.05)
    ea, eb = EMAModel(a, decay=0.999).bind(a), EMAModel(b, decay=0.999)
    ref = EMAOracle([p.detach().cpu() for p in a], decay=0.999)
    for step in range(1, 5):
        gen = torch.Generator().manual_seed(50 + step)
        for pa, pb, g in zip(a, b, gs):
            gr = (torch.randn(g.shape, generator=gen) * 1e-2).to(dtype).to(cuda)
            pa.grad, pb.grad = gr.clone(), gr.clone()
        oa.clip_and_step(max_norm=1.0, ema=ea)
        ob.clip_and_step(max_norm=1.0)
        eb.step(b)
        ref.step([p.detach().cpu().requires_grad_(True) for p in a])
        for pa, pb in zip(a, b):
            assert torch.equal(pa, pb)
        for sa, sb, sr in zip(ea.shadow_params, eb.shadow_params, ref.shadow_params):
            assert torch.equal(sa, sb)
            assert torch.equal(sa.cpu(), sr)
    assert ea.optimization_step == eb.optimization_step == 4


def test_optimizer_state_dict_reload_uses_the_loaded_moments(cuda):
    """step -> state_dict -> load_state_dict -> step: the device tables must follow the re-allocated moment
    tensors (the table cache is keyed on parameter AND moment pointers), and the result equals an optimizer that
    never went through the reload."""
    from yat_b200.optim import FusedClipAdamW
    ps, gs = _make(torch.bfloat16, 31, 1e-2)
    a = [torch.nn.Parameter(p.to(cuda)) for p in ps]
    b = [torch.nn.Parameter(p.to(cuda)) for p in ps]
    oa, ob = FusedClipAdamW(a, lr=1e-2), FusedClipAdamW(b, lr=1e-2)

    def grads(step):
        gen = torch.Generator().manual_seed(70 + step)
        return [(torch.randn(g.shape, generator=gen) * 1e-2).to(torch.bfloat16).to(cuda) for g in gs]

    for step in (1, 2):
        for pa, pb, g in zip(a, b, grads(step)):
            pa.grad, pb.grad = g.clone(), g.clone()
        oa.step(max_norm=1.0), ob.step(max_norm=1.0)
    import io
    buf = io.BytesIO()
    torch.save(ob.state_dict(), buf)                        # a checkpoint: the reload allocates every state tensor anew
    buf.seek(0)
    sd = torch.load(buf, map_location="cpu")
    old = [ob.state[p]["exp_avg"] for p in b]
    ob.load_state_dict(sd)                                  # replaces every state tensor
    assert all(ob.state[p]["exp_avg"].data_ptr() != o.data_ptr() for p, o in zip(b, old))
    for o in old:
        o.fill_(float("nan"))                               # a stale table would read (and spread) these
    del old
    for step in (3, 4):
        for pa, pb, g in zip(a, b, grads(step)):
            pa.grad, pb.grad = g.clone(), g.clone()
        oa.step(max_norm=1.0), ob.step(max_norm=1.0)
    for pa, pb in zip(a, b):
        assert torch.equal(pa, pb)
        assert torch.equal(oa.state[pa]["exp_avg"], ob.state[pb]["exp_avg"])
        assert torch.equal(oa.state[pa]["exp_avg_sq"], ob.state[pb]["exp_avg_sq"])


def test_mixed_step_counts_in_one_group_are_refused(cuda):
    from yat_b200.optim import FusedClipAdamW
    a = torch.nn.Parameter(torch.zeros(64, device=cuda, dtype=torch.bfloat16))
    b = torch.nn.Parameter(torch.zeros(64, device=cuda, dtype=torch.bfloat16))
    opt = FusedClipAdamW([a, b], lr=1e-2)
    a.grad, b.grad = torch.ones_like(a), torch.ones_like(b)
    opt.step()
    opt.state[b]["step"] += 5
    a.grad, b.grad = torch.ones_like(a), torch.ones_like(b)
    with pytest.raises(ValueError, match="different steps"):
        opt.step()
