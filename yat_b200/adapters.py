"""LoRA / LoKr adapters for the trainer (the PEFT wrap of common/trainer.py:212-241), as thin PyTorch glue.

Upstream calls ``peft.get_peft_model(model, LoraConfig | LoKrConfig(r, alpha, target_modules))``; peft is not part
of this image.  An adapter here is a small sub-module ``<target>.adapter`` whose contribution is added in
*activation space*, like peft's unmerged forward - merging ``W0 + dW`` into a bf16 weight would round small
updates away (a bf16 ulp of a 0.02-sized weight is 1e-4):

    LoRA   y += scale * (x A^T) B^T                 A [r, in], B [out, r], scale = alpha / r, B = 0 at init
    LoKr   y += scale * kron(W1, W2) x              out = ol * ok, in = im * in_ (factors closest to sqrt),
           = scale * W1 (X W2^T), X = x as [im, in_]  W1 [ol, im], W2 = W2a W2b ([ok, r] [r, in_]; kept full when
                                                    r is not smaller than half its sides), W2b (or W2) = 0 at init

The base weights stay frozen, ordinary ``weight`` tensors, so the hand-written kernels and the token-major GEMM
path of GLUMBConv see the same layout as in a full fine-tune; modules reached as ``module(x)`` get the term from a
forward hook, ``GLUMBConv.forward_tokens`` (which reads ``.weight`` directly) adds it through ``adapter_term``.
Initialisation: LoRA as peft (A kaiming-uniform, B zero).  LoKr keeps ``w1`` kaiming-uniform and zero-initialises the
``w2`` side; peft's ``LoKrLayer.reset_adapter_parameters`` zeroes ``w1`` and draws ``w2`` - either way the adapter starts at
exactly 0, the two differ only in which factor receives the first non-zero gradient (stated deviation, unpinned:
peft is not installed).
``adapters_disabled(model)`` runs the base weights only: the preservation term's frozen prediction
(README.md:81-82; SURVEY.md §8 A7: "for LoRA/LoKr the same weights with adapters disabled") needs no second copy of
the model.  Parity with peft is unpinned (the package is absent); the formulas are the published LoRA
(arXiv 2106.09685) / LyCORIS LoKr ones and ``tests/test_host_logic.py`` pins them against explicit fp64 evaluation.
"""
import contextlib
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


def _factor(n):
    """n = a * b with a <= b as close to sqrt(n) as possible (LyCORIS ``factorization(n, -1)``)."""
    a = int(math.isqrt(n))
    while n % a:
        a -= 1
    return a, n // a


class _Adapter(nn.Module):
    enabled = True

    def delta_weight(self):
        """[out, in] weight-space form of the adapter (tests, merging for export)."""
        raise NotImplementedError


class LoRAAdapter(_Adapter):
    def __init__(self, out_features, in_features, rank, alpha, dtype, device):
        super().__init__()
        self.scale = float(alpha) / float(rank)
        self.lora_A = nn.Parameter(torch.empty((rank, in_features), dtype=dtype, device=device))
        self.lora_B = nn.Parameter(torch.zeros((out_features, rank), dtype=dtype, device=device))
        nn.init.kaiming_uniform_(self.lora_A, a=math.sqrt(5))

    def forward(self, x):
        return F.linear(F.linear(x, self.lora_A), self.lora_B) * self.scale

    def delta_weight(self):
        return self.scale * (self.lora_B @ self.lora_A)


class _ScaledKron(torch.autograd.Function):
    """scale * w1[a, b] * w2[c, d] as [a, c, b, d]; the backward contracts the dense gradient with the other factor as
    two einsums (permute + matmul) - autograd's own path through the broadcast product reduces 5-25 M elements with
    strided torch reductions per adapted module, which was ~10 % of the K5 step."""

    @staticmethod
    def forward(ctx, w1, w2, scale):
        ctx.save_for_backward(w1, w2)
        ctx.scale = scale
        return scale * (w1[:, None, :, None] * w2[None, :, None, :])

    @staticmethod
    def backward(ctx, g):
        w1, w2 = ctx.saved_tensors
        g = g * ctx.scale if ctx.scale != 1.0 else g
        dw1 = torch.einsum("acbd,cd->ab", g, w2) if ctx.needs_input_grad[0] else None
        dw2 = torch.einsum("acbd,ab->cd", g, w1) if ctx.needs_input_grad[1] else None
        return dw1, dw2, None


class LoKrAdapter(_Adapter):
    def __init__(self, out_features, in_features, rank, alpha, dtype, device):
        super().__init__()
        self.scale = float(alpha) / float(rank)
        ol, ok = _factor(out_features)
        im, in_ = _factor(in_features)
        self.shape = (ol, ok, im, in_)
        self.lokr_w1 = nn.Parameter(torch.empty((ol, im), dtype=dtype, device=device))
        nn.init.kaiming_uniform_(self.lokr_w1, a=math.sqrt(5))
        if rank < max(ok, in_) / 2:
            self.lokr_w2_a = nn.Parameter(torch.empty((ok, rank), dtype=dtype, device=device))
            self.lokr_w2_b = nn.Parameter(torch.zeros((rank, in_), dtype=dtype, device=device))
            nn.init.kaiming_uniform_(self.lokr_w2_a, a=math.sqrt(5))
            self.lokr_w2 = None
        else:
            self.lokr_w2 = nn.Parameter(torch.zeros((ok, in_), dtype=dtype, device=device))
            self.lokr_w2_a = self.lokr_w2_b = None

    def _w2(self):
        return self.lokr_w2 if self.lokr_w2 is not None else self.lokr_w2_a @ self.lokr_w2_b

    # above this many tokens the Kronecker factor is expanded to a dense [out, in] matrix (a few MB, rebuilt per call)
    # and applied as ONE tensor-core GEMM: the factored form is 16 Ki tiny batched matmuls plus reshapes, which made
    # the SANA-1.6B K5 step 2x slower than the full fine-tune on B200 (350 vs 162 ms).  Same product, other rounding order.
    dense_tokens = 1024

    def forward(self, x):
        ol, ok, im, in_ = self.shape
        lead = x.shape[:-1]
        if x.is_cuda and x.numel() // x.shape[-1] >= self.dense_tokens:
            return F.linear(x, self.delta_weight().to(x.dtype))
        t = F.linear(x.reshape(*lead, im, in_), self._w2())               # [.., im, ok]  = X W2^T
        y = torch.matmul(self.lokr_w1, t)                                  # [.., ol, ok]  = W1 (X W2^T)
        return y.reshape(*lead, ol * ok) * self.scale

    def delta_weight(self):
        ol, ok, im, in_ = self.shape
        # kron(w1, w2)[a * ok + c, b * in_ + d] = w1[a, b] * w2[c, d]
        return _ScaledKron.apply(self.lokr_w1, self._w2(), self.scale).reshape(ol * ok, im * in_)


ALGOS = {"lora": LoRAAdapter, "lokr": LoKrAdapter}


def _matches(name, targets):
    return any(name == t or name.endswith("." + t) for t in targets)


def _hook(mod, inputs, output):
    ad = mod.adapter
    if not ad.enabled:
        return output
    x = inputs[0]
    if isinstance(mod, nn.Conv2d):                                         # NCHW module path of a 1x1 convolution
        return output + ad(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)
    return output + ad(x)


def adapter_term(mod, x):
    """Adapter contribution of ``mod`` for token-major ``x`` ([.., in]), or None - for code that applies
    ``mod.weight`` itself instead of calling the module (GLUMBConv.forward_tokens)."""
    ad = getattr(mod, "adapter", None)
    if ad is None or not ad.enabled:
        return None
    return ad(x)


def apply_adapters(model, algo, rank, alpha, target_modules):
    """Freeze ``model`` and attach an adapter to every Linear / 1x1 Conv2d whose qualified name ends with one of
    ``target_modules`` (peft's suffix rule; a string is split on whitespace like the YAML reader leaves it).
    Returns the list of adapted module names."""
    if algo not in ALGOS:
        raise ValueError(f"lora_algo {algo!r} is not supported here (have: {sorted(ALGOS)})")
    targets = target_modules.split() if isinstance(target_modules, str) else list(target_modules)
    for p in model.parameters():
        p.requires_grad_(False)
    done = []
    for name, mod in list(model.named_modules()):
        if not _matches(name, targets):
            continue
        if isinstance(mod, nn.Linear):
            out_f, in_f = mod.weight.shape
        elif isinstance(mod, nn.Conv2d) and mod.kernel_size == (1, 1) and mod.groups == 1:
            out_f, in_f = mod.weight.shape[0], mod.weight.shape[1]
        else:
            raise ValueError(f"{name}: adapters attach to Linear / 1x1 Conv2d, not {type(mod).__name__} "
                             "(upstream README.md:16 excludes conv_depth for the same reason)")
        mod.adapter = ALGOS[algo](out_f, in_f, int(rank), float(alpha if alpha is not None else rank), mod.weight.dtype,
                                  mod.weight.device)
        mod.register_forward_hook(_hook)
        done.append(name)
    if not done:
        raise ValueError(f"no module matches lora_target_modules={targets}")
    return done


def adapter_modules(model):
    return [m for m in model.modules() if isinstance(m, _Adapter)]


def has_adapters(model):
    return bool(adapter_modules(model))


@contextlib.contextmanager
def adapters_disabled(model):
    mods = adapter_modules(model)
    for m in mods:
        m.enabled = False
    try:
        yield
    finally:
        for m in mods:
            m.enabled = True


def adapter_state_dict(model):
    """Only the adapter tensors (what ``save_pretrained`` of a PeftModel stores)."""
    return {k: v for k, v in model.state_dict().items() if ".adapter." in k}


ADAPTER_CONFIG_NAME = "adapter_config.json"
ADAPTER_WEIGHTS_NAME = "adapter_model.safetensors"


def save_adapters(model, path, algo, rank, alpha, target_modules):
    """Adapter-only checkpoint in peft's file layout (``adapter_config.json`` + ``adapter_model.safetensors``):
    what ``PeftModel.save_pretrained`` writes upstream (common/trainer.py:295-296 on a peft-wrapped model).  The
    frozen base weights are not part of it, so the directory reloads with ``load_adapters`` on a fresh base model."""
    import json
    import os
    from safetensors.torch import save_file
    os.makedirs(path, exist_ok=True)
    targets = target_modules.split() if isinstance(target_modules, str) else list(target_modules)
    with open(os.path.join(path, ADAPTER_CONFIG_NAME), "w") as f:
        json.dump({"peft_type": algo.upper(), "r": int(rank), "alpha": float(alpha if alpha is not None else rank),
                   "target_modules": targets}, f, indent=2)
    save_file({k: v.detach().contiguous().cpu() for k, v in adapter_state_dict(model).items()},
              os.path.join(path, ADAPTER_WEIGHTS_NAME))


def load_adapters(model, path):
    """Attach adapters described by ``path/adapter_config.json`` to ``model`` and load their tensors."""
    import json
    import os
    from safetensors.torch import load_file
    with open(os.path.join(path, ADAPTER_CONFIG_NAME)) as f:
        cfg = json.load(f)
    names = apply_adapters(model, cfg["peft_type"].lower(), cfg["r"], cfg["alpha"], cfg["target_modules"])
    sd = load_file(os.path.join(path, ADAPTER_WEIGHTS_NAME))
    own = adapter_state_dict(model)
    if set(sd) != set(own):
        raise ValueError(f"adapter checkpoint keys do not match the model: {sorted(set(sd) ^ set(own))[:4]} ...")
    for k, v in sd.items():
        own[k].copy_(v.to(own[k].device, own[k].dtype))
    return names
