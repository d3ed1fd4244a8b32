"""Scheduler tables (oracle; test infrastructure only).

Restates the two diffusers schedulers the reference instantiates:
  * ``FlowMatchEulerDiscreteScheduler`` — train_sana.py:41, train_sd35.py:49
  * ``DDPMScheduler``                    — train_pixart_sigma.py:37
diffusers is not vendored in the reference tree (requirements.txt:7, unpinned); the
formulas below are its published ``__init__`` arithmetic, anchored on the reference's
call sites ``self.scheduler.timesteps[indices]`` (train_sana.py:193) and
``self.scheduler.sigmas`` (train_sana.py:196).
"""
import numpy as np
import torch


def flow_match_tables(num_train_timesteps: int = 1000, shift: float = 1.0):
    """(timesteps fp32 [T], sigmas fp32 [T]) of FlowMatchEulerDiscreteScheduler.__init__.

    timesteps = linspace(1, T, T)[::-1]; sigma = t / T;
    sigma <- shift*sigma / (1 + (shift-1)*sigma); timesteps = sigma * T.
    """
    t = np.linspace(1, num_train_timesteps, num_train_timesteps, dtype=np.float32)[::-1].copy()
    t = torch.from_numpy(t).to(dtype=torch.float32)
    sigmas = t / num_train_timesteps
    sigmas = shift * sigmas / (1 + (shift - 1) * sigmas)
    timesteps = sigmas * num_train_timesteps
    return timesteps, sigmas


def ddpm_tables(num_train_timesteps: int = 1000, beta_start: float = 1e-4, beta_end: float = 0.02):
    """(timesteps int64 [T] descending, alphas_cumprod fp32 [T]) of DDPMScheduler.__init__
    with beta_schedule="linear" (PixArt-Sigma's scheduler config)."""
    betas = torch.linspace(beta_start, beta_end, num_train_timesteps, dtype=torch.float32)
    alphas = 1.0 - betas
    alphas_cumprod = torch.cumprod(alphas, dim=0)
    timesteps = torch.from_numpy(np.arange(0, num_train_timesteps)[::-1].copy())
    return timesteps, alphas_cumprod


def get_sigmas(timesteps, schedule_timesteps, sigmas, n_dim=4, dtype=torch.float32):
    """train_sana.py:195-204 — index by float equality, cast table to ``dtype`` first."""
    sig = sigmas.to(dtype=dtype)
    step_indices = [(schedule_timesteps == t).nonzero().item() for t in timesteps]
    sigma = sig[step_indices].flatten()
    while len(sigma.shape) < n_dim:
        sigma = sigma.unsqueeze(-1)
    return sigma
