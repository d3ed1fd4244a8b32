"""Scheduler tables, device-resident.

Host-side mirror of the two diffusers schedulers the reference's hooks instantiate
(train_sana.py:41, train_sd35.py:49 — FlowMatchEulerDiscreteScheduler; train_pixart_sigma.py:37 —
DDPMScheduler): same attribute names (``config.num_train_timesteps``, ``timesteps``, ``sigmas``,
``alphas_cumprod``, ``scale_noise``, ``add_noise``) so hook code reads the same.  The tables are
built once (the reference re-uploads them and does B ``.item()`` lookups every step,
train_sana.py:195-204) and live on the device; lookups happen inside the fused noising kernel.
"""
from types import SimpleNamespace

import numpy as np
import torch


class FlowMatchEulerDiscreteScheduler:
    def __init__(self, num_train_timesteps: int = 1000, shift: float = 1.0):
        self.config = SimpleNamespace(num_train_timesteps=num_train_timesteps, shift=shift)
        t = np.linspace(1, num_train_timesteps, num_train_timesteps, dtype=np.float32)[::-1].copy()
        t = torch.from_numpy(t).to(dtype=torch.float32)
        sigmas = t / num_train_timesteps
        sigmas = shift * sigmas / (1 + (shift - 1) * sigmas)
        self.timesteps = sigmas * num_train_timesteps     # fp32 [T], CPU (as in diffusers)
        self.sigmas = sigmas                              # fp32 [T], CPU
        self._dev = {}

    def device_tables(self, device):
        """(sigmas, timesteps) as fp32 device tensors, uploaded once per device."""
        key = str(device)
        if key not in self._dev:
            self._dev[key] = (self.sigmas.to(device).contiguous(), self.timesteps.to(device).contiguous())
        return self._dev[key]

    def scale_noise(self, sample, timestep_indices, noise):
        """x_t for given table indices through the fused kernel (train_sd35.py:185 call shape,
        except that indices — not float timesteps — select the row: no equality search)."""
        from .noise import flowmatch_noise
        u = (timestep_indices.to(torch.float32) + 0.5) / self.config.num_train_timesteps
        return flowmatch_noise(sample, self, noise=noise, u=u.to(sample.device), want_target=False).x_t


class DDPMScheduler:
    def __init__(self, num_train_timesteps: int = 1000, beta_start: float = 1e-4, beta_end: float = 0.02):
        self.config = SimpleNamespace(num_train_timesteps=num_train_timesteps, beta_start=beta_start,
                                      beta_end=beta_end, beta_schedule="linear")
        self.betas = torch.linspace(beta_start, beta_end, num_train_timesteps, dtype=torch.float32)
        self.alphas = 1.0 - self.betas
        self.alphas_cumprod = torch.cumprod(self.alphas, dim=0)
        self.timesteps = torch.from_numpy(np.arange(0, num_train_timesteps)[::-1].copy())
        self._dev = {}

    def device_tables(self, device):
        key = str(device)
        if key not in self._dev:
            self._dev[key] = self.alphas_cumprod.to(device).contiguous()
        return self._dev[key]

    def add_noise(self, original_samples, noise, timesteps):
        """diffusers signature (train_pixart_sigma.py:176); ``timesteps`` are int64 DDPM steps."""
        from .noise import ddpm_noise
        T = self.config.num_train_timesteps
        idx = (T - 1) - timesteps.to(torch.int64)
        u = (idx.to(torch.float32) + 0.5) / T
        return ddpm_noise(original_samples, self, noise=noise, u=u.to(original_samples.device)).x_t
