"""Clip + AdamW as two persistent kernels (host side of ``yat_grad_clip_coef`` / ``yat_adamw_step``).

Replaces ``accelerator.clip_grad_norm_(params, 1.0)`` + ``torch.optim.AdamW.step()`` of the step loop
(common/trainer.py:245-248,347-348).  ``FusedClipAdamW`` is a ``torch.optim.Optimizer``
(same ``param_groups`` / ``state_dict`` layout as ``torch.optim.AdamW``: ``step``, ``exp_avg``,
``exp_avg_sq`` per parameter, so LambdaLR and checkpoints keep working).  Every parameter of a group is cut
into fixed-size chunks listed in device tables built once; per step only the gradient pointers are
refreshed (a few KB, pinned staging, no host sync), the total gradient norm and the clip factor stay on
the device, and the factor is applied to the gradient inside the update.  No CPU fallback.
"""
import torch

from . import _lib

_DTYPES = {torch.bfloat16: 0, torch.float32: 1}
_RING = 4   # pinned staging buffers for the gradient-pointer table (the host may run ahead of the GPU)


class _Group:
    """Device tables of one (param group, dtype) set of parameters."""

    def __init__(self, params, device):
        lib = _lib.lib()
        self.params = params
        self.dtype = params[0].dtype
        chunk = lib.yat_optim_chunk_elems()
        rows, chunks = [], []
        for i, p in enumerate(params):
            chunks += [(i, c) for c in range((p.numel() + chunk - 1) // chunk)]
        self.n_chunks = len(chunks)
        self.chunks = torch.tensor(chunks, dtype=torch.int32).to(device)
        self.tab = None                                   # built when the state exists
        self.gtab = torch.zeros(len(params), dtype=torch.int64, device=device)
        self.ring = [torch.zeros(len(params), dtype=torch.int64).pin_memory() for _ in range(_RING)]
        self.events = [None] * _RING
        self.slot = 0
        self.ws = torch.zeros(lib.yat_optim_workspace_bytes(self.n_chunks), dtype=torch.uint8, device=device)
        self.norm_coef = torch.zeros(2, dtype=torch.float32, device=device)

    def build_table(self, state):
        rows = []
        for p in self.params:
            st = state[p]
            rows.append([p.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(), p.numel()])
        self.tab = torch.tensor(rows, dtype=torch.int64).to(self.gtab.device)
        self.tab_key = self.key_of(state)
        self.etab = self.etab_key = None

    def key_of(self, state):
        """Every pointer the device table caches: Optimizer.load_state_dict() swaps the moment tensors while the
        parameter pointers stay, so the parameters alone do not identify a valid table."""
        return tuple((p.data_ptr(), state[p]["exp_avg"].data_ptr(), state[p]["exp_avg_sq"].data_ptr())
                     for p in self.params)

    def ema_table(self, shadows):
        key = tuple(s.data_ptr() for s in shadows)
        if self.etab is None or self.etab_key != key:
            for p, s in zip(self.params, shadows):
                if s.dtype != p.dtype or s.shape != p.shape or not s.is_contiguous() or s.device != p.device:
                    raise ValueError("FusedClipAdamW: EMA shadow tensors must match their parameters (dtype, shape, device)")
            self.etab = torch.tensor(key, dtype=torch.int64).to(self.gtab.device)
            self.etab_key = key
        return self.etab

    def upload_grads(self):
        """Gradient pointers of this step -> device (async from a pinned ring slot)."""
        k = self.slot
        self.slot = (k + 1) % _RING
        if self.events[k] is not None:
            self.events[k].synchronize()                  # that copy finished long ago in steady state
        host = self.ring[k]
        host.copy_(torch.tensor([p.grad.data_ptr() for p in self.params], dtype=torch.int64))
        self.gtab.copy_(host, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self.events[k] = ev


class FusedClipAdamW(torch.optim.Optimizer):
    """AdamW with the global-norm gradient clip folded in.  ``step(max_norm=...)`` = clip + update;
    ``step()`` alone = plain AdamW (torch.optim.AdamW semantics, decoupled weight decay)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self._groups = {}
        self.last_norm = None     # device scalar: total gradient norm of the last clipped step

    def _sets(self):
        """[(group dict, _Group)] over the parameters that have a gradient, split by dtype."""
        out = []
        for gi, group in enumerate(self.param_groups):
            by_dtype = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise ValueError("FusedClipAdamW runs on CUDA parameters only (no CPU fallback)")
                if p.dtype not in _DTYPES or p.grad.dtype != p.dtype:
                    raise ValueError(f"FusedClipAdamW: unsupported dtype {p.dtype} / grad {p.grad.dtype}")
                if not p.is_contiguous() or not p.grad.is_contiguous():
                    raise ValueError("FusedClipAdamW: parameters and gradients must be contiguous")
                by_dtype.setdefault(p.dtype, []).append(p)
            for dtype, ps in by_dtype.items():
                key = (gi, dtype, tuple(id(p) for p in ps))
                g = self._groups.get(key)
                if g is None:
                    g = self._groups[key] = _Group(ps, ps[0].device)
                for p in ps:
                    st = self.state[p]
                    if len(st) == 0:
                        st["step"] = torch.zeros((), dtype=torch.float32)
                        st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                        st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                if g.tab is None or g.tab_key != g.key_of(self.state):
                    g.build_table(self.state)
                out.append((group, g))
        return out

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._groups = {}          # the device tables point at the moment tensors that were just replaced

    def __setstate__(self, state):
        super().__setstate__(state)
        self._groups = {}

    @torch.no_grad()
    def step(self, closure=None, max_norm=None, ema=None):
        """``ema``: a bound ``yat_b200.ema.EMAModel``; its shadow tensors are updated in the same pass as the
        weights (common/trainer.py:350-351 ``ema_model.step(parameters)`` right after ``optimizer.step()``)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.lib()
        sets = self._sets()
        if not sets:
            return loss
        stream = _lib.current_stream()
        for _, g in sets:
            g.upload_grads()
        coef = None
        if max_norm is not None:
            if len(sets) != 1:
                raise ValueError("FusedClipAdamW: the fused clip needs all clipped parameters in one group and dtype")
            g = sets[0][1]
            with _lib.timed("grad_clip_coef"):
                rc = lib.yat_grad_clip_coef(_lib.ptr(g.tab), _lib.ptr(g.gtab), _lib.ptr(g.chunks), g.n_chunks,
                                            _DTYPES[g.dtype], float(max_norm), _lib.ptr(g.norm_coef), _lib.ptr(g.ws),
                                            stream)
            _lib.check(rc, "yat_grad_clip_coef")
            coef = g.norm_coef[1:]
            self.last_norm = g.norm_coef[0]
        omd = ema.advance() if ema is not None else 0.0
        for group, g in sets:
            steps = set()
            for p in g.params:
                st = self.state[p]["step"]
                st += 1
                steps.add(int(st))
            if len(steps) != 1:
                # the bias corrections are launch constants: a tensor that first got a gradient later would be
                # updated with the wrong 1 - beta^t
                raise ValueError(f"FusedClipAdamW: parameters of one group are at different steps {sorted(steps)}; "
                                 "give parameters that start receiving gradients later their own param group")
            step = steps.pop()
            b1, b2 = group["betas"]
            etab = g.ema_table(ema.shadow_of(g.params)) if ema is not None else None
            with _lib.timed("adamw_step"):
                rc = lib.yat_adamw_ema_step(_lib.ptr(g.tab), _lib.ptr(g.gtab), _lib.ptr(g.chunks), g.n_chunks,
                                            _DTYPES[g.dtype], _lib.ptr(coef), float(group["lr"]), float(b1), float(b2),
                                            float(group["eps"]), float(group["weight_decay"]), step,
                                            _lib.ptr(etab), float(omd), stream)
            _lib.check(rc, "yat_adamw_ema_step")
        if ema is not None:
            done = {id(p) for _, g in sets for p in g.params}
            rest = [(s, p) for s, p in zip(ema.shadow_params, ema._params()) if id(p) not in done]
            for s, p in rest:
                if not p.requires_grad:
                    ema.copy_if_changed(s, p)                   # EMAModel.step: frozen tensors are copied
                elif p.grad is None:
                    raise ValueError("FusedClipAdamW: a trainable parameter tracked by the EMA received no gradient; "
                                     "call EMAModel.step(parameters) instead of the fused path")
        return loss

    def clip_and_step(self, max_norm=1.0, ema=None):
        """``accelerator.clip_grad_norm_(params, max_norm); optimizer.step(); ema_model.step(params)`` in two launches."""
        return self.step(max_norm=max_norm, ema=ema)
