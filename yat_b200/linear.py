"""Biased Linear layers of the SANA block with the bias gradient from an own column-sum kernel.

torch autograd forms ``grad_bias = grad_output.sum(0)`` with a strided ``reduce_kernel`` (≈ 2 TB/s on
[16384, 2240] bf16: 151 launches, 5.3 ms of a SANA-1.6B step in ``profiles/r01d_launches.md``).  ``Linear`` keeps
``nn.Linear``'s parameters, state-dict keys and forward GEMM (cuBLAS, bias in the epilogue) and replaces only that
reduction by ``yat_colsum_bf16`` (lanes along the columns, fp32 partials, fixed-order finish).  Anything the kernel
does not cover (CPU tensors in the structure tests, fp32 modules, no bias, no gradient wanted) is the plain
``F.linear`` autograd path - that is the same arithmetic, not a fallback of a product kernel.
Reference: the ``to_q / to_k / to_v / to_out`` projections of diffusers ``Attention`` inside the block restated at
utils/patch_sana_attention_layers.py:72-115.
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


def colsum(a):
    """[rows, C] bf16 CUDA, contiguous -> [C] bf16 (fp32 accumulation)."""
    if not a.is_cuda or a.dtype != torch.bfloat16 or a.dim() != 2 or not a.is_contiguous():
        raise ValueError("colsum wants a contiguous [rows, C] bf16 CUDA matrix (no CPU fallback)")
    rows, C = a.shape
    out = torch.empty((C,), dtype=torch.bfloat16, device=a.device)
    ws = torch.empty(_lib.lib().yat_colsum_workspace_bytes(rows, C), dtype=torch.uint8, device=a.device)
    with _lib.timed("colsum"):
        rc = _lib.lib().yat_colsum_bf16(_lib.ptr(a), rows, C, _lib.ptr(out), _lib.ptr(ws), _lib.current_stream())
    _lib.check(rc, "yat_colsum_bf16")
    return out


class _LinearBias(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        ctx.save_for_backward(x, weight)
        return F.linear(x, weight, bias)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        need = ctx.needs_input_grad
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = (dy2 @ weight).view_as(x) if need[0] else None
        dw = dy2.t() @ x.reshape(-1, x.shape[-1]) if need[1] else None
        db = colsum(dy2) if need[2] else None
        return dx, dw, db


def linear_bias(x, weight, bias):
    if (bias is not None and x.is_cuda and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16
            and bias.dtype == torch.bfloat16 and weight.shape[0] % 8 == 0 and bias.requires_grad
            and torch.is_grad_enabled()):
        return _LinearBias.apply(x, weight, bias)
    return F.linear(x, weight, bias)


class Linear(nn.Linear):
    def forward(self, x):
        return linear_bias(x, self.weight, self.bias)
