"""Projection GEMM on tcgen05 (host side of ``yat_gemm_bf16_tn``): ``c = a @ w.T (+ bias)``.

The ``x @ W^T`` form of every projection in the SANA block (diffusers ``Attention.to_q/to_k/to_v/to_out`` and the
token-major 1x1 convolutions of ``GLUMBConv``; utils/patch_sana_attention_layers.py:84-113), SURVEY.md §8(f) N1.
``a`` is ``[..., K]`` bf16 (rows may be a strided view with unit stride along K), ``w`` ``[N, K]`` bf16, ``bias``
``[N]`` bf16 or None.  No CPU fallback.
"""
import torch

from . import _lib


def gemm_bf16_tn(a, w, bias=None, tile_n=0, out=None):
    if not (a.is_cuda and w.is_cuda):
        raise ValueError("yat_b200 ops run on CUDA tensors only (no CPU fallback)")
    if a.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or (bias is not None and bias.dtype != torch.bfloat16):
        raise ValueError("gemm_bf16_tn expects bf16 operands")
    if w.dim() != 2 or a.shape[-1] != w.shape[1]:
        raise ValueError(f"shapes {tuple(a.shape)} x {tuple(w.shape)}^T do not contract")
    lead = a.shape[:-1]
    K = a.shape[-1]
    a2 = a.reshape(-1, K)
    if a2.stride(1) != 1 or a2.stride(0) % 8 or a2.data_ptr() % 16:
        a2 = a2.contiguous()
    if w.stride(1) != 1 or w.stride(0) % 8 or w.data_ptr() % 16:
        w = w.contiguous()
    M, N = a2.shape[0], w.shape[0]
    if bias is not None:
        bias = bias.contiguous()
        if bias.numel() != N:
            raise ValueError(f"bias has {bias.numel()} elements, expected {N}")
    ldc = (N + 15) // 16 * 16
    if out is None:
        buf = torch.empty((M, ldc), dtype=torch.bfloat16, device=a.device)
    else:
        buf = out.view(M, -1)
        if buf.stride(1) != 1 or buf.stride(0) % 16 or buf.stride(0) < N or buf.data_ptr() % 32:
            raise ValueError("out must be [M, >= N] with a row pitch that is a multiple of 16 elements, 32-byte aligned")
        ldc = buf.stride(0)
    with _lib.timed("gemm_bf16_tn"):
        rc = _lib.lib().yat_gemm_bf16_tn(_lib.ptr(a2), _lib.ptr(w), _lib.ptr(bias), _lib.ptr(buf), M, N, K,
                                         a2.stride(0), w.stride(0), ldc, int(tile_n), _lib.current_stream())
    _lib.check(rc, "yat_gemm_bf16_tn")
    c = buf[:, :N] if buf.shape[1] != N else buf
    return c.reshape(*lead, N) if c.is_contiguous() else c.contiguous().reshape(*lead, N)
