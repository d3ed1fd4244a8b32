"""One optimisation step of the reference's SANA trainer on the CPU (oracle; test infrastructure only).

Restates, in the reference's order and with the reference's ATen ops:
  common/trainer.py:325       generator = torch.Generator()           (fresh, unseeded, every step)
  train_sana.py:167-179       pad each [L_i, E] embedding to 512 tokens, int64 mask, stack
  train_sana.py:183-193       noise, then u, from the same generator; idx; t
  train_sana.py:195-207       get_sigmas (float-equality lookup), x_t
  train_sana.py:210-215       transformer forward (oracle.sana_model.sana_forward, checkpointed
                              like train_sana.py:63)
  train_sana.py:217-218       target, fp32 MSE
  common/trainer.py:344-356   backward, clip_grad_norm_(1.0), optimizer.step(), zero_grad()
Used by the parity tests (tiny config) and as the timed CPU baseline / ``--impl reference`` arm of
bench.py (SANA-1.6B shapes, bounded batch).
"""
import torch
import torch.nn.functional as F

from . import losses, noising, sampling, schedule
from .sana_model import sana_forward


def pad_embeddings_reference(embeddings, max_tokens=512):
    """train_sana.py:167-179: python loop of F.pad + ones/zeros mask, then stack."""
    embeds, masks = [], []
    for e in embeddings:
        n = min(e.shape[0], max_tokens)
        embeds.append(F.pad(e[:n], (0, 0, 0, max_tokens - n)))
        m = torch.zeros(max_tokens, dtype=torch.int64)
        m[:n] = 1
        masks.append(m)
    return torch.stack(embeds), torch.stack(masks)


def random_state_dict(model_meta, dtype=torch.float32, seed=0):
    """Random-init HF-layout state dict with the shapes of ``model_meta`` (a module built on the
    meta device — shapes only, no product compute): N(0, 0.02) matrices, zero biases, unit norms."""
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for name, p in model_meta.state_dict().items():
        if name.endswith("bias"):
            t = torch.zeros(p.shape, dtype=dtype)
        elif name.endswith("caption_norm.weight"):
            t = torch.ones(p.shape, dtype=dtype)
        elif "scale_shift_table" in name:
            t = (torch.randn(p.shape, generator=g) / p.shape[-1] ** 0.5).to(dtype)
        else:
            t = torch.empty(p.shape, dtype=torch.float32).normal_(0.0, 0.02, generator=g).to(dtype)
        sd[name] = t.requires_grad_(True)
    return sd


def train_step(sd, cfg, optimizer, latents, embeddings, shift=1.0, generator=None, max_tokens=512,
               gradient_checkpointing=True, compute_dtype=None):
    """latents [B,C,h,w] (training dtype), embeddings: list of [L_i, E].  Returns the loss (float).

    ``compute_dtype``: dtype the model runs in (default: the latents' dtype, like the reference's
    bf16 pipeline; pass torch.float32 for BASELINE.json configs[0] "fp32 on CPU")."""
    if generator is None:
        generator = torch.Generator()                                                  # trainer.py:325
    enc, mask = pad_embeddings_reference(embeddings, max_tokens)
    noise, u, idx = sampling.sana_step_draws(latents.shape, latents.shape[0], latents.dtype, generator)
    tt, ss = schedule.flow_match_tables(1000, shift)
    x_t, target, t, _ = noising.flow_noise_sana(latents, noise, idx, tt, ss)
    cd = compute_dtype or latents.dtype
    pred = sana_forward(sd, cfg, x_t.to(cd), enc.to(cd), t, mask, gradient_checkpointing=gradient_checkpointing)
    loss = losses.mse_fp32(pred, target)
    loss.backward()
    params = [p for p in sd.values() if p.grad is not None]
    torch.nn.utils.clip_grad_norm_(params, 1.0)                                        # trainer.py:347
    optimizer.step()
    optimizer.zero_grad(set_to_none=True)
    return float(loss.detach())
