"""Clip + AdamW step of the training loop (oracle; test infrastructure only).

Restates ``accelerator.clip_grad_norm_(model.parameters(), max_norm=1.0); optimizer.step()``
(common/trainer.py:347-348) with ``torch.optim.AdamW(params, lr, weight_decay)`` (common/trainer.py:245-248,
default betas (0.9, 0.999), eps 1e-8):

* ``reference_step`` runs exactly those two torch calls on CPU copies (the reference's own code path:
  foreach ops that round to the storage dtype after every op);
* ``exact_step`` evaluates the same update in float64 from the same stored inputs and rounds each stored
  tensor once - the arithmetic the device kernel follows in fp32 (yat_b200/csrc/optim.cu).
"""
import torch


def reference_step(params, grads, state, lr, weight_decay, max_norm=1.0, betas=(0.9, 0.999), eps=1e-8):
    """params/grads: lists of CPU tensors (any float dtype).  state: None or the optimizer returned by a
    previous call.  Returns (new params, total_norm, optimizer)."""
    ps = [torch.nn.Parameter(p.clone()) for p in params] if state is None else state.param_groups[0]["params"]
    opt = state or torch.optim.AdamW(ps, lr=lr, weight_decay=weight_decay, betas=betas, eps=eps, foreach=False)
    for p, g in zip(ps, grads):
        p.grad = g.clone()
    norm = torch.nn.utils.clip_grad_norm_(ps, max_norm)
    opt.step()
    return [p.detach().clone() for p in ps], norm, opt


def exact_step(params, grads, exp_avg, exp_avg_sq, step, lr, weight_decay, max_norm=1.0, betas=(0.9, 0.999),
               eps=1e-8):
    """float64 evaluation on the stored values, one rounding per stored tensor.  ``step`` is the 1-based
    step count of this update.  Returns (p, m, v lists in the input dtype, total_norm float, clip coef float)."""
    b1, b2 = betas
    total = sum(float((g.double() ** 2).sum()) for g in grads) ** 0.5
    coef = 1.0 if max_norm is None else min(1.0, max_norm / (total + 1e-6))
    coef32 = float(torch.tensor(coef, dtype=torch.float32))
    bc1, bc2 = 1.0 - b1 ** step, 1.0 - b2 ** step
    out_p, out_m, out_v = [], [], []
    for p, g, m, v in zip(params, grads, exp_avg, exp_avg_sq):
        dt = p.dtype
        g64 = (g.float() * coef32).to(dt).double()            # clip_grad_norm_ leaves the gradient in its dtype
        p64 = p.double() * (1.0 - lr * weight_decay)
        m64 = m.double() + (1.0 - b1) * (g64 - m.double())
        v64 = b2 * v.double() + (1.0 - b2) * g64 * g64
        p64 = p64 - (lr / bc1) * m64 / (v64.sqrt() / bc2 ** 0.5 + eps)
        out_p.append(p64.to(dt))
        out_m.append(m64.to(dt))
        out_v.append(v64.to(dt))
    return out_p, out_m, out_v, total, coef
