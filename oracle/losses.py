"""Denoising losses (oracle; test infrastructure only)."""
import torch


def mse_fp32(pred, target):
    """train_sana.py:182,218 — ``MSELoss()(pred.float(), target.float())`` (fp32 scalar)."""
    return torch.nn.MSELoss()(pred.float(), target.float())


def mse_lowp(pred, target):
    """train_sd35.py:179,193 and train_pixart_sigma.py:170,184 —
    ``MSELoss()(pred.to(noise.dtype), target)`` evaluated in the training dtype (bf16)."""
    return torch.nn.MSELoss()(pred.to(dtype=target.dtype), target)


def pixart_crop(pred_2c, noise):
    """train_pixart_sigma.py:182-183 — learned-sigma head: first half of the channels is the
    epsilon prediction; the target is the noise cropped to that shape."""
    noise_pred = pred_2c.chunk(2, 1)[0]
    target = noise[: noise_pred.shape[0], : noise_pred.shape[1], : noise_pred.shape[2], : noise_pred.shape[3]]
    return noise_pred, target


def denoise_loss(pred, x0, noise, mode="velocity", pred_frozen=None, rho=0.0, weights=None,
                 lowp=False):
    """Generalised loss the fused kernel implements; reduces to the reference for
    ``weights=None, pred_frozen=None``.

    mode "velocity": target = noise - x0 in the latent dtype (train_sana.py:217);
    mode "eps":      target = noise                  (train_pixart_sigma.py:183).
    weights [B]:     L_noise = (1/B) sum_b w_b * mean_chw (pred-target)^2  (w=1 -> global mean).
    pred_frozen:     use_preservation term (README.md:81-82; not implemented upstream — our
                     stated formula): L = L_noise + rho * MSE(pred.float(), pred_frozen.float()).
    lowp:            bf16 loss semantics of mse_lowp (result rounded to the training dtype).
    Returns (total, noise_term, preservation_term).
    """
    C = x0.shape[1]
    p = pred[:, :C]
    target = (noise - x0) if mode == "velocity" else noise
    if weights is None and not lowp:
        l_noise = mse_fp32(p, target)
    elif weights is None and lowp:
        l_noise = mse_lowp(p, target)
    else:
        d = (p.float() - target.float()) ** 2
        per_sample = d.flatten(1).mean(dim=1)
        l_noise = (per_sample * weights.float()).mean()
        if lowp:
            l_noise = l_noise.to(target.dtype)
    l_pres = torch.zeros((), dtype=l_noise.dtype)
    total = l_noise
    if pred_frozen is not None:
        l_pres = torch.nn.MSELoss()(p.float(), pred_frozen[:, :C].float())
        total = l_noise + rho * l_pres.to(l_noise.dtype)
    return total, l_noise, l_pres


def sd3_loss_weighting(weighting_scheme, sigmas):
    """diffusers.training_utils.compute_loss_weighting_for_sd3 restated (the SD3 / SANA trainers of diffusers weight the
    per-sample flow-matching MSE by it; the YAT hooks train_sd35.py:179-193 / train_sana.py:217-218 use no weighting,
    i.e. ``none``): sigma_sqrt -> sigma^-2, cosmap -> 2 / (pi (1 - 2 sigma + 2 sigma^2)), anything else -> 1."""
    import math
    import torch
    sigmas = sigmas.float()
    if weighting_scheme == "sigma_sqrt":
        return sigmas ** -2.0
    if weighting_scheme == "cosmap":
        return 2.0 / (math.pi * (1.0 - 2.0 * sigmas + 2.0 * sigmas ** 2))
    return torch.ones_like(sigmas)
