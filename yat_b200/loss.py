"""Fused denoising loss (host side of ``yat_loss_fused``).

One kernel computes the scalar loss *and* dL/dpred; the autograd node just hands the stored
gradient back.  Covers train_sana.py:217-218 (fp32 velocity MSE), train_sd35.py:192-193 (bf16
velocity MSE), train_pixart_sigma.py:182-184 (bf16 eps MSE on the first half of the channels),
per-sample weights, and the use_preservation term (README.md:81-82).
"""
import torch

from . import _lib

VELOCITY, EPS = 0, 1

_workspaces = {}


def _workspace(device):
    key = str(device)
    if key not in _workspaces:
        n = _lib.lib().yat_loss_workspace_bytes()
        _workspaces[key] = torch.zeros(n, dtype=torch.uint8, device=device)  # zeroed once; kernel keeps it so
    return _workspaces[key]


class _FusedDenoiseLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, x0, noise, pred_frozen, weights, rho, gscale, target_mode, lowp, seed, offset,
                trust_gscale):
        if not pred.is_cuda:
            raise ValueError("yat_b200 ops run on CUDA tensors only (no CPU fallback)")
        if pred.dtype != torch.bfloat16:
            raise ValueError("fused loss expects a bf16 prediction")
        ref = x0 if x0 is not None else noise
        B, C = ref.shape[0], ref.shape[1]
        hw = ref[0, 0].numel()
        if pred.shape[0] != B or pred.shape[1] < C or pred[0, 0].numel() != hw:
            raise ValueError(f"pred shape {tuple(pred.shape)} incompatible with latents {tuple(ref.shape)}")
        pred_c = pred.contiguous()
        x0_c = None if x0 is None else x0.contiguous()
        noise_c = None if noise is None else noise.contiguous()
        pf_c = None if pred_frozen is None else pred_frozen.contiguous()
        w_c = None if weights is None else weights.to(device=pred.device, dtype=torch.float32).contiguous()
        need_grad = pred.requires_grad
        dpred = torch.empty_like(pred_c) if need_grad else None
        out = torch.empty(3, dtype=torch.float32, device=pred.device)
        with _lib.timed("loss_fused"):
            rc = _lib.lib().yat_loss_fused(
                _lib.ptr(pred_c), _lib.ptr(x0_c), _lib.ptr(noise_c), _lib.ptr(pf_c), _lib.ptr(w_c), float(rho),
                float(gscale), int(target_mode), 1 if lowp else 0, seed, offset, B, C, pred.shape[1], hw,
                _lib.ptr(out), _lib.ptr(dpred), _lib.ptr(_workspace(pred.device)), _lib.current_stream())
        _lib.check(rc, "yat_loss_fused")
        ctx.save_for_backward(dpred)
        ctx.gscale = float(gscale)
        ctx.trust = bool(trust_gscale)
        ctx.mark_non_differentiable(out)
        loss = out[0].to(torch.bfloat16) if lowp else out[0].clone()
        return loss, out

    @staticmethod
    def backward(ctx, grad_loss, _grad_terms):
        (dpred,) = ctx.saved_tensors
        if dpred is None:
            return (None,) * 12
        if not ctx.trust:
            # dpred was computed for an upstream gradient of `gscale`; rescale to the actual one
            dpred = dpred * (grad_loss.to(torch.float32) / ctx.gscale).to(dpred.dtype)
        return (dpred,) + (None,) * 11


def fused_denoise_loss(pred, x0, noise=None, *, pred_frozen=None, weights=None, rho=0.0, gscale=1.0,
                       target_mode=VELOCITY, lowp=False, seed=0, offset=0, trust_gscale=False,
                       return_terms=False):
    """Scalar loss with grad wrt ``pred``.

    noise=None regenerates eps from ``(seed, offset)`` (must match the noising call).
    gscale is the upstream gradient the stored dpred is pre-multiplied by (accelerate's
    1/gradient_accumulation_steps); with ``trust_gscale`` the backward is a pure hand-off
    (true single pass), otherwise it is rescaled by the actual upstream gradient.
    """
    loss, terms = _FusedDenoiseLoss.apply(pred, x0, noise, pred_frozen, weights, rho, gscale, target_mode, lowp,
                                          seed, offset, trust_gscale)
    return (loss, terms) if return_terms else loss


def sd3_loss_weights(weighting_scheme, sigmas):
    """Per-sample loss weights w_b of the SD3 weighting schemes from the sigmas the noising kernel left on the device
    (no host sync): ``sigma_sqrt`` -> sigma^-2, ``cosmap`` -> 2 / (pi (1 - 2 sigma + 2 sigma^2)); ``none`` / None -> None
    (w = 1: what train_sd35.py:179-193 and train_sana.py:217-218 compute).  Feeds ``fused_denoise_loss(weights=)``."""
    if weighting_scheme in (None, "none", "logit_normal"):
        return None
    sigmas = sigmas.float().reshape(-1)
    if weighting_scheme == "sigma_sqrt":
        return sigmas ** -2.0
    if weighting_scheme == "cosmap":
        return 2.0 / (3.141592653589793 * (1.0 - 2.0 * sigmas + 2.0 * sigmas * sigmas))
    raise ValueError(f"unknown weighting_scheme {weighting_scheme!r} (none, sigma_sqrt, cosmap)")
