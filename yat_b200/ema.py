"""EMA of the trained weights: host side of ``yat_ema_step`` / ``yat_adamw_ema_step``.

Mirror of the slice of ``diffusers.training_utils.EMAModel`` the reference trainer uses
(common/trainer.py:264-268 ``EMAModel(self.model.parameters(), decay=0.999)``; :350-351
``ema_model.step(self.model.parameters())`` after every optimiser step; :374-377 mean-reduce of
``shadow_params`` over ranks at validation; :381-383,399-400 ``store`` / ``copy_to`` / ``restore`` around
validation + save).  diffusers is not installed, so the class is restated from its published
behaviour (parity unpinned by the reference):

    optimization_step += 1
    decay = get_decay(optimization_step)      # min(decay, (1 + s) / (10 + s)) with s = step - 1; 0 while s <= 0
    s_param.sub_((1 - decay) * (s_param - param))   per trainable tensor;  s_param.copy_(param) for frozen ones

Shadow tensors are clones of the parameters (same dtype: bf16 for the reference's bf16 models).  The arithmetic
runs in ``optim.cu`` over the optimiser's chunk tables, with every op rounded to the storage dtype like the three
tensor ops upstream.  ``FusedClipAdamW.step(..., ema=...)`` folds the update into the AdamW pass
(+4 B/param of HBM traffic); ``EMAModel.step(parameters)`` alone is one extra streaming kernel.  No CPU fallback.
"""
import torch

from . import _lib

_DTYPES = {torch.bfloat16: 0, torch.float32: 1}


class EMAModel:
    def __init__(self, parameters, decay: float = 0.9999, min_decay: float = 0.0, update_after_step: int = 0,
                 use_ema_warmup: bool = False, inv_gamma: float = 1.0, power: float = 2 / 3):
        parameters = list(parameters)
        self.shadow_params = [p.clone().detach() for p in parameters]
        self.temp_stored_params = None
        self.decay = decay
        self.min_decay = min_decay
        self.update_after_step = update_after_step
        self.use_ema_warmup = use_ema_warmup
        self.inv_gamma = inv_gamma
        self.power = power
        self.optimization_step = 0
        self.cur_decay_value = None
        self._tables = {}

    # ---- diffusers EMAModel.get_decay ----
    def get_decay(self, optimization_step: int) -> float:
        step = max(0, optimization_step - self.update_after_step - 1)
        if step <= 0:
            return 0.0
        if self.use_ema_warmup:
            cur = 1 - (1 + step / self.inv_gamma) ** -self.power
        else:
            cur = (1 + step) / (10 + step)
        return max(min(cur, self.decay), self.min_decay)

    def advance(self) -> float:
        """Count one optimiser step; returns ``1 - decay`` for it (the value both kernels take)."""
        self.optimization_step += 1
        decay = self.get_decay(self.optimization_step)
        self.cur_decay_value = decay
        return 1.0 - decay

    def shadow_of(self, params):
        """Shadow tensors aligned with ``params`` (used by the fused optimiser path); identity by position."""
        if not hasattr(self, "_index"):
            raise RuntimeError("EMAModel.bind(parameters) must be called before the fused path is used")
        return [self.shadow_params[self._index[id(p)]] for p in params]

    def bind(self, parameters):
        """Remember which shadow belongs to which parameter object (order of construction)."""
        parameters = list(parameters)
        if len(parameters) != len(self.shadow_params):
            raise ValueError("EMAModel.bind: parameter list does not match the shadow list")
        self._index = {id(p): i for i, p in enumerate(parameters)}
        self._bound = parameters
        return self

    def _params(self):
        return self._bound

    def copy_if_changed(self, shadow, param):
        """``s_param.copy_(param)`` of EMAModel.step for a frozen tensor, skipped while the tensor has not been
        written since the last copy (the shadow is a clone; with LoRA / LoKr the frozen base is ~all of the model)."""
        seen = self.__dict__.setdefault("_versions", {})
        if seen.get(id(param)) != param._version:
            shadow.copy_(param)
            seen[id(param)] = param._version

    @torch.no_grad()
    def step(self, parameters):
        """``ema_model.step(model.parameters())`` (common/trainer.py:350-351): one kernel per dtype."""
        parameters = list(parameters)
        if len(parameters) != len(self.shadow_params):
            raise ValueError("EMAModel.step: parameter list does not match the shadow list")
        omd = self.advance()
        lib = _lib.lib()
        by_dtype = {}
        for s, p in zip(self.shadow_params, parameters):
            if not p.is_cuda:
                raise ValueError("EMAModel.step runs on CUDA parameters only (no CPU fallback)")
            if not p.requires_grad:
                s.copy_(p)                                   # frozen tensors track the weights exactly
                continue
            if p.dtype not in _DTYPES or s.dtype != p.dtype or not p.is_contiguous() or not s.is_contiguous():
                raise ValueError(f"EMAModel.step: unsupported tensor ({p.dtype}, contiguous={p.is_contiguous()})")
            by_dtype.setdefault(p.dtype, []).append((p, s))
        chunk = lib.yat_optim_chunk_elems()
        for dtype, pairs in by_dtype.items():
            key = (dtype, tuple(p.data_ptr() for p, _ in pairs), tuple(s.data_ptr() for _, s in pairs))
            tabs = self._tables.get(dtype)
            if tabs is None or tabs[0] != key:
                rows = [[p.data_ptr(), s.data_ptr(), 0, p.numel()] for p, s in pairs]
                chunks = [(i, c) for i, (p, _) in enumerate(pairs) for c in range((p.numel() + chunk - 1) // chunk)]
                dev = pairs[0][0].device
                tabs = (key, torch.tensor(rows, dtype=torch.int64).to(dev),
                        torch.tensor(chunks, dtype=torch.int32).to(dev), len(chunks))
                self._tables[dtype] = tabs
            with _lib.timed("ema_step"):
                rc = lib.yat_ema_step(_lib.ptr(tabs[1]), _lib.ptr(tabs[2]), tabs[3], _DTYPES[dtype], float(omd),
                                      _lib.current_stream())
            _lib.check(rc, "yat_ema_step")

    # ---- validation-time helpers (plain tensor copies, not on the per-step path) ----
    def to(self, device=None, dtype=None):
        self.shadow_params = [s.to(device=device, dtype=dtype) if s.is_floating_point() else s.to(device=device)
                              for s in self.shadow_params]
        self._tables = {}
        return self

    @torch.no_grad()
    def copy_to(self, parameters):
        for s, p in zip(self.shadow_params, list(parameters)):
            p.data.copy_(s.to(p.device).data)

    @torch.no_grad()
    def store(self, parameters):
        self.temp_stored_params = [p.detach().clone() for p in parameters]

    @torch.no_grad()
    def restore(self, parameters):
        if self.temp_stored_params is None:
            raise RuntimeError("This ExponentialMovingAverage has no `store()`ed weights to `restore()`")
        for c, p in zip(self.temp_stored_params, list(parameters)):
            p.data.copy_(c.data)
        self.temp_stored_params = None

    @torch.no_grad()
    def reduce_mean(self, accelerator):
        """common/trainer.py:374-377: every rank averages its shadow tensors before validation."""
        for s in self.shadow_params:
            s.data = accelerator.reduce(s.data, reduction="mean")
        self._tables = {}

    def state_dict(self):
        return {"decay": self.decay, "min_decay": self.min_decay, "optimization_step": self.optimization_step,
                "update_after_step": self.update_after_step, "use_ema_warmup": self.use_ema_warmup,
                "inv_gamma": self.inv_gamma, "power": self.power, "shadow_params": self.shadow_params}

    def load_state_dict(self, sd):
        for k in ("decay", "min_decay", "optimization_step", "update_after_step", "use_ema_warmup", "inv_gamma", "power"):
            if k in sd:
                setattr(self, k, sd[k])
        if sd.get("shadow_params") is not None:
            self.shadow_params = [t.clone() for t in sd["shadow_params"]]
        self._tables = {}
