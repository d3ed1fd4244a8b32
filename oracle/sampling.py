"""Timestep and noise sampling (oracle; test infrastructure only)."""
import torch


def logit_normal_u(batch_size, generator=None, logit_mean=0.0, logit_std=1.0):
    """diffusers ``compute_density_for_timestep_sampling('logit_normal', ...)`` as called at
    train_sana.py:185-191 and common/trainer.py:98: ``u = sigmoid(normal(mean, std, (B,)))``
    drawn on the CPU from ``generator`` (``mode_scale`` is ignored for this scheme)."""
    u = torch.normal(mean=logit_mean, std=logit_std, size=(batch_size,), device="cpu", generator=generator)
    return torch.nn.functional.sigmoid(u)


def timestep_indices(u, num_train_timesteps=1000, clamp=True):
    """train_sana.py:192 / common/trainer.py:99: ``(u * T).long()``.

    ``clamp`` caps the index at T-1: fp32 sigmoid saturates to exactly 1.0 for z > ~16.6,
    which makes the reference index out of range (latent upstream bug); the device path
    clamps and so does the oracle (documented in DESIGN.md)."""
    idx = (u * num_train_timesteps).long()
    if clamp:
        idx = idx.clamp_(max=num_train_timesteps - 1)
    return idx


def randn_like_reference(shape, dtype, generator=None):
    """diffusers ``randn_tensor`` with a CPU generator (train_sana.py:183): draw on the CPU
    in the target dtype, then copy to the device."""
    return torch.randn(shape, generator=generator, device="cpu", dtype=dtype)


def sana_step_draws(shape, batch_size, dtype=torch.bfloat16, generator=None, T=1000):
    """Order of draws in train_sana.py:183-192: noise first, then u, from the same generator."""
    if generator is None:
        generator = torch.Generator()  # common/trainer.py:325 — fresh, unseeded, every step
    noise = randn_like_reference(shape, dtype, generator)
    u = logit_normal_u(batch_size, generator)
    idx = timestep_indices(u, T)
    return noise, u, idx
