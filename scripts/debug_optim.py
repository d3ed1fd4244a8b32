import sys
import torch
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from oracle import optimizer as oopt
from yat_b200.optim import FusedClipAdamW
from test_optim_gpu import SIZES, _make
dev = torch.device("cuda")
dtype = torch.bfloat16
lr, wd = 1e-2, 0.05
ps, gs = _make(dtype, 5, 1e-4)
dev_ps = [torch.nn.Parameter(p.to(dev)) for p in ps]
opt = FusedClipAdamW(dev_ps, lr=lr, weight_decay=wd)
for step in range(1, 4):
    gen = torch.Generator().manual_seed(100 + step)
    grads = [(torch.randn(g.shape, generator=gen) * 1e-4 * (1.0 + 0.25 * step)).to(dtype) for g in gs]
    for p, g in zip(dev_ps, grads):
        p.grad = g.to(dev)
    cur = [p.detach().cpu().clone() for p in dev_ps]
    m = [opt.state[p]["exp_avg"].cpu().clone() if len(opt.state[p]) else torch.zeros_like(q) for p, q in zip(dev_ps, ps)]
    v = [opt.state[p]["exp_avg_sq"].cpu().clone() if len(opt.state[p]) else torch.zeros_like(q) for p, q in zip(dev_ps, ps)]
    m_in, v_in, p_in = [t.clone() for t in m], [t.clone() for t in v], [t.clone() for t in cur]
    opt.step(max_norm=1.0)
    cur, m, v, norm, coef = oopt.exact_step(cur, grads, m, v, step, lr, wd, max_norm=1.0)
    print("step", step, "norm", opt.last_norm.item(), norm, coef)
    for ti in range(len(ps)):
        for name, got_t, ref_t in (("p", dev_ps[ti].detach(), cur[ti]), ("m", opt.state[dev_ps[ti]]["exp_avg"], m[ti]),
                                   ("v", opt.state[dev_ps[ti]]["exp_avg_sq"], v[ti])):
            got, ref = got_t.float().cpu().flatten(), ref_t.float().flatten()
            d = (got - ref).abs()
            bad = (d > 2.0 ** -7 * ref.abs().clamp_min(1e-30) * 1.01).nonzero().flatten().tolist()
            if int((d > 0).sum()):
                print("  tensor", ti, name, "mismatch", int((d > 0).sum()), "bad", len(bad))
            for i in bad[:4]:
                print("     ", i, "got", repr(got[i].item()), "ref", repr(ref[i].item()), "| p_in", p_in[ti].flatten()[i].item(), "g", grads[ti].flatten()[i].item(),
                      "m_in", m_in[ti].flatten()[i].item(), "v_in", v_in[ti].flatten()[i].item())
