"""Fused adaLN-single modulate / gate / residual ops of the SANA block (host side of
``yat_ln_modulate_*`` and ``yat_gate_residual_*``).

The reference block (utils/patch_sana_attention_layers.py:82-115) computes
``norm(x) * (1 + scale) + shift`` and ``x + gate * y`` as LayerNorm plus broadcast elementwise
launches; here each is one CUDA kernel per direction.  ``scale``/``shift``/``gate`` are the
``[B, 1, C]`` chunks of ``scale_shift_table[None] + timestep.reshape(B, 6, -1)`` and are read in
place through a batch stride (no copies).  No CPU fallback.
"""
import torch

from . import _lib


def _mod_view(t, B, C, name):
    """[B, 1, C] / [B, C] bf16 with unit stride along C -> (tensor, batch stride in elements)."""
    if t.dim() == 3:
        if t.shape[1] != 1:
            raise ValueError(f"{name} must be [B, 1, C], got {tuple(t.shape)}")
        t = t[:, 0]
    if tuple(t.shape) != (B, C):
        raise ValueError(f"{name} must have shape [{B}, {C}], got {tuple(t.shape)}")
    if t.dtype != torch.bfloat16:
        raise ValueError(f"{name} must be bf16")
    if t.stride(1) != 1 or t.stride(0) % 8 or t.data_ptr() % 16 or (B > 1 and t.stride(0) < C):
        t = t.contiguous()
    return t, (t.stride(0) if B > 1 else C)


def _check_x(x, name):
    if not x.is_cuda:
        raise ValueError("yat_b200 ops run on CUDA tensors only (no CPU fallback)")
    if x.dtype != torch.bfloat16 or x.dim() != 3:
        raise ValueError(f"{name} must be a [B, N, C] bf16 tensor")


def _workspace(B, N, C, device):
    return torch.empty(_lib.lib().yat_block_workspace_bytes(B, N, C), dtype=torch.uint8, device=device)


class _LnModulate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, shift, eps):
        _check_x(x, "x")
        B, N, C = x.shape
        x = x.contiguous()
        sc, sc_stride = _mod_view(scale, B, C, "scale")
        sh, sh_stride = _mod_view(shift, B, C, "shift")
        if sh_stride != sc_stride:
            sc, sh = sc.contiguous(), sh.contiguous()
            sc_stride = C
        need_grad = any(ctx.needs_input_grad[:3])
        y = torch.empty_like(x)
        mean = torch.empty(B * N, dtype=torch.float32, device=x.device) if need_grad else None
        rstd = torch.empty(B * N, dtype=torch.float32, device=x.device) if need_grad else None
        with _lib.timed("ln_modulate_fwd"):
            rc = _lib.lib().yat_ln_modulate_fwd(_lib.ptr(x), _lib.ptr(sc), _lib.ptr(sh), sc_stride, B, N, C, eps,
                                                _lib.ptr(y), _lib.ptr(mean), _lib.ptr(rstd), _lib.current_stream())
        _lib.check(rc, "yat_ln_modulate_fwd")
        if need_grad:
            ctx.save_for_backward(x, sc, mean, rstd)
            ctx.sc_stride = sc_stride
            ctx.mod_shapes = (scale.shape, shift.shape)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, sc, mean, rstd = ctx.saved_tensors
        B, N, C = x.shape
        dy = dy.contiguous()
        dx = torch.empty_like(x)
        dscale = torch.empty((B, C), dtype=torch.bfloat16, device=x.device)
        dshift = torch.empty((B, C), dtype=torch.bfloat16, device=x.device)
        ws = _workspace(B, N, C, x.device)
        with _lib.timed("ln_modulate_bwd"):
            rc = _lib.lib().yat_ln_modulate_bwd(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(sc), ctx.sc_stride, _lib.ptr(mean),
                                                _lib.ptr(rstd), B, N, C, _lib.ptr(dx), _lib.ptr(dscale),
                                                _lib.ptr(dshift), _lib.ptr(ws), _lib.current_stream())
        _lib.check(rc, "yat_ln_modulate_bwd")
        s_shape, h_shape = ctx.mod_shapes
        return dx, dscale.view(s_shape), dshift.view(h_shape), None


class _LnModulateRes(torch.autograd.Function):
    """(y, h) = (ln_modulate(h), h) with h = x [+ x2]: the normalised branch AND the residual branch of a block
    sub-layer leave one autograd node, so the two gradients that flow back into h are added inside the LN backward
    kernel (``dres``) instead of by an autograd accumulation pass, and the attn2 residual add (x + x2,
    utils/patch_sana_attention_layers.py:103) is formed inside the LN forward kernel."""

    @staticmethod
    def forward(ctx, x, x2, scale, shift, eps):
        _check_x(x, "x")
        B, N, C = x.shape
        x = x.contiguous()
        if x2 is not None:
            _check_x(x2, "x2")
            if x2.shape != x.shape:
                raise ValueError(f"x {tuple(x.shape)} and x2 {tuple(x2.shape)} differ")
            x2 = x2.contiguous()
        sc, sc_stride = _mod_view(scale, B, C, "scale")
        sh, sh_stride = _mod_view(shift, B, C, "shift")
        if sh_stride != sc_stride:
            sc, sh = sc.contiguous(), sh.contiguous()
            sc_stride = C
        need_grad = any(ctx.needs_input_grad[:4])
        y = torch.empty_like(x)
        h = torch.empty_like(x) if x2 is not None else x
        mean = torch.empty(B * N, dtype=torch.float32, device=x.device) if need_grad else None
        rstd = torch.empty(B * N, dtype=torch.float32, device=x.device) if need_grad else None
        with _lib.timed("ln_modulate_fwd"):
            rc = _lib.lib().yat_add_ln_modulate_fwd(_lib.ptr(x), _lib.ptr(x2), _lib.ptr(sc), _lib.ptr(sh), sc_stride,
                                                    B, N, C, eps, _lib.ptr(h) if x2 is not None else None,
                                                    _lib.ptr(y), _lib.ptr(mean), _lib.ptr(rstd), _lib.current_stream())
        _lib.check(rc, "yat_add_ln_modulate_fwd")
        if need_grad:
            ctx.save_for_backward(h, sc, mean, rstd)
            ctx.sc_stride = sc_stride
            ctx.mod_shapes = (scale.shape, shift.shape)
            ctx.has_x2 = x2 is not None
        if x2 is None:
            h = x.view_as(x)        # a fresh alias of the input: its gradient arrives at THIS node's backward
        return y, h

    @staticmethod
    def backward(ctx, dy, dh):
        h, sc, mean, rstd = ctx.saved_tensors
        B, N, C = h.shape
        if dy is None:              # only the residual branch was used downstream
            return dh, (dh if ctx.has_x2 else None), None, None, None
        dy = dy.contiguous()
        dh = None if dh is None else dh.contiguous()
        dx = torch.empty_like(h)
        dscale = torch.empty((B, C), dtype=torch.bfloat16, device=h.device)
        dshift = torch.empty((B, C), dtype=torch.bfloat16, device=h.device)
        ws = _workspace(B, N, C, h.device)
        with _lib.timed("ln_modulate_bwd"):
            rc = _lib.lib().yat_ln_modulate_bwd_res(_lib.ptr(h), _lib.ptr(dy), _lib.ptr(dh), _lib.ptr(sc), ctx.sc_stride,
                                                    _lib.ptr(mean), _lib.ptr(rstd), B, N, C, _lib.ptr(dx),
                                                    _lib.ptr(dscale), _lib.ptr(dshift), _lib.ptr(ws),
                                                    _lib.current_stream())
        _lib.check(rc, "yat_ln_modulate_bwd_res")
        s_shape, h_shape = ctx.mod_shapes
        return dx, (dx if ctx.has_x2 else None), dscale.view(s_shape), dshift.view(h_shape), None


class _GateResidual(torch.autograd.Function):
    @staticmethod
    def forward(ctx, res, y, gate):
        _check_x(res, "res")
        _check_x(y, "y")
        if res.shape != y.shape:
            raise ValueError(f"res {tuple(res.shape)} and y {tuple(y.shape)} differ")
        B, N, C = res.shape
        res, y = res.contiguous(), y.contiguous()
        g, g_stride = _mod_view(gate, B, C, "gate")
        out = torch.empty_like(res)
        with _lib.timed("gate_residual_fwd"):
            rc = _lib.lib().yat_gate_residual_fwd(_lib.ptr(res), _lib.ptr(y), _lib.ptr(g), g_stride, B, N, C,
                                                  _lib.ptr(out), _lib.current_stream())
        _lib.check(rc, "yat_gate_residual_fwd")
        if any(ctx.needs_input_grad):
            ctx.save_for_backward(y, g)
            ctx.g_stride = g_stride
            ctx.gate_shape = gate.shape
        return out

    @staticmethod
    def backward(ctx, dout):
        y, g = ctx.saved_tensors
        B, N, C = y.shape
        dout = dout.contiguous()
        dy = torch.empty_like(y)
        dgate = torch.empty((B, C), dtype=torch.bfloat16, device=y.device)
        ws = _workspace(B, N, C, y.device)
        with _lib.timed("gate_residual_bwd"):
            rc = _lib.lib().yat_gate_residual_bwd(_lib.ptr(dout), _lib.ptr(y), _lib.ptr(g), ctx.g_stride, B, N, C,
                                                  _lib.ptr(dy), _lib.ptr(dgate), _lib.ptr(ws), _lib.current_stream())
        _lib.check(rc, "yat_gate_residual_bwd")
        return dout, dy, dgate.view(ctx.gate_shape)


def ln_modulate(x, scale, shift, eps=1e-6):
    """LayerNorm(x, no affine, eps) * (1 + scale) + shift;  x [B,N,C] bf16, scale/shift [B,1,C] bf16."""
    return _LnModulate.apply(x, scale, shift, float(eps))


def ln_modulate_res(x, scale, shift, eps=1e-6, add=None):
    """(ln_modulate(h), h) with h = x (``add`` None) or bf16(x + add): use the returned h for the residual branch, so
    that both gradients of h are combined inside the LN backward kernel."""
    return _LnModulateRes.apply(x, add, scale, shift, float(eps))


def gate_residual(res, y, gate):
    """res + gate * y;  res, y [B,N,C] bf16, gate [B,1,C] bf16."""
    return _GateResidual.apply(res, y, gate)
