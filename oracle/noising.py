"""Noising step (oracle; test infrastructure only).

Runs the reference's own ATen ops on the CPU, in the reference's dtype and order, so the
bf16 rounding chain (four roundings for x_t) is reproduced by construction.
"""
import torch

from . import schedule as _sched


def flow_noise_sana(latents, noise, idx, timesteps_tab, sigmas_tab):
    """train_sana.py:193-207,217.

    latents/noise: [B,C,h,w] in the training dtype (bf16, or fp32 for config K1).
    Returns (x_t, target, t, sigma) with t fp32 [B] and sigma in the latent dtype [B,1,1,1].
    """
    t = timesteps_tab[idx]
    sigmas = _sched.get_sigmas(t, timesteps_tab, sigmas_tab, latents.ndim, dtype=latents.dtype)
    x_t = (1.0 - sigmas) * latents + sigmas * noise          # train_sana.py:207
    target = noise - latents                                  # train_sana.py:217
    return x_t, target, t, sigmas


def flow_noise_sd3(latents, noise, idx, timesteps_tab, sigmas_tab):
    """train_sd35.py:183-185,192 — ``scheduler.scale_noise(latents, timesteps, noise)``:
    ``sigma * noise + (1.0 - sigma) * sample`` with sigma = sigmas.to(sample.dtype)[index]."""
    t = timesteps_tab[idx]
    sigmas = _sched.get_sigmas(t, timesteps_tab, sigmas_tab, latents.ndim, dtype=latents.dtype)
    x_t = sigmas * noise + (1.0 - sigmas) * latents
    target = noise - latents
    return x_t, target, t, sigmas


def ddpm_noise(latents, noise, idx, timesteps_tab, alphas_cumprod):
    """train_pixart_sigma.py:173-176,183 — ``DDPMScheduler.add_noise``; target is the noise.

    alphas_cumprod is cast to the sample dtype *before* ``** 0.5`` (diffusers behaviour)."""
    t = timesteps_tab[idx]                                      # int64, = T-1-idx
    ac = alphas_cumprod.to(dtype=latents.dtype)
    sqrt_alpha_prod = ac[t] ** 0.5
    sqrt_alpha_prod = sqrt_alpha_prod.flatten()
    while len(sqrt_alpha_prod.shape) < len(latents.shape):
        sqrt_alpha_prod = sqrt_alpha_prod.unsqueeze(-1)
    sqrt_one_minus_alpha_prod = (1 - ac[t]) ** 0.5
    sqrt_one_minus_alpha_prod = sqrt_one_minus_alpha_prod.flatten()
    while len(sqrt_one_minus_alpha_prod.shape) < len(latents.shape):
        sqrt_one_minus_alpha_prod = sqrt_one_minus_alpha_prod.unsqueeze(-1)
    x_t = sqrt_alpha_prod * latents + sqrt_one_minus_alpha_prod * noise
    return x_t, noise, t, (sqrt_alpha_prod, sqrt_one_minus_alpha_prod)
