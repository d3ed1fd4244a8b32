"""REPA wrapper model and the fused cosine-alignment loss.

Host-side mirror of common/repa.py: ``RepaConfig`` (:11-41), ``RepaMLP`` (:43-57), ``RepaModule``
(:59-80), ``RepaModel`` patch/unpatch + save/load layout (:82-212) keep their names, fields and
on-disk files (``repa_config.json`` + ``repa_model_weights.bin``).  ``calculate_loss`` is where
this build differs on purpose: the reference's is a no-grad token-mean MSE placeholder
(:137-154, never added to the loss, common/trainer.py:340-341); here it is the REPA cosine
alignment with a gradient, computed by ``yat_repa_cos``.
"""
import json
import os
from dataclasses import dataclass, field
from typing import Optional, Union

import torch

from . import _lib


@dataclass
class RepaConfig:
    target_modules: Optional[Union[list, str]] = field(default=None)
    target_shape: int = field(default=768)
    hidden_shape: list = field(default=None)


class RepaMLP(torch.nn.Module):
    def __init__(self, in_size, hidden_size, out_size):
        super().__init__()
        self.fc1 = torch.nn.Linear(in_size, hidden_size)
        self.relu = torch.nn.ReLU()
        self.fc2 = torch.nn.Linear(hidden_size, out_size)

    def forward(self, x):
        return self.fc2(self.relu(self.fc1(x)))


class RepaModule(torch.nn.Module):
    def __init__(self, base_module: torch.nn.Module, in_size: int, out_size: int):
        super().__init__()
        self.zs = None
        self.base_module = base_module
        self.proj_mlp = RepaMLP(in_size, out_size, out_size)

    def forward(self, *args, **kwargs):
        hidden_states = self.base_module(*args, **kwargs)
        self.zs = self.proj_mlp(hidden_states)
        return hidden_states


class _RepaCos(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z, y, grid_hw, teacher_hw):
        if not z.is_cuda:
            raise ValueError("yat_b200 ops run on CUDA tensors only (no CPU fallback)")
        if z.dtype != torch.bfloat16:
            raise ValueError("projected student tokens must be bf16")
        B, N, E = z.shape
        gh, gw = grid_hw
        th, tw = teacher_hw
        if N != gh * gw or tuple(y.shape) != (B, th * tw, E):
            raise ValueError(f"z {tuple(z.shape)} / y {tuple(y.shape)} do not match grids {grid_hw}->{teacher_hw}")
        if y.dtype not in (torch.bfloat16, torch.float32):
            raise ValueError("teacher tokens must be bf16 or fp32")
        z_c, y_c = z.contiguous(), y.contiguous()
        lib = _lib.lib()
        ws = torch.zeros(lib.yat_repa_workspace_bytes(B, th, tw, E), dtype=torch.uint8, device=z.device)
        loss = torch.empty(1, dtype=torch.float32, device=z.device)
        dz = torch.empty_like(z_c) if z.requires_grad else None
        with _lib.timed("repa_cos"):
            rc = lib.yat_repa_cos(_lib.ptr(z_c), _lib.ptr(y_c), 0 if y.dtype == torch.bfloat16 else 1, B, gh, gw, th, tw,
                                  E, 1.0, _lib.ptr(loss), _lib.ptr(dz), _lib.ptr(ws), _lib.current_stream())
        _lib.check(rc, "yat_repa_cos")
        ctx.save_for_backward(dz)
        return loss[0]

    @staticmethod
    def backward(ctx, grad):
        (dz,) = ctx.saved_tensors
        if dz is None:
            return None, None, None, None
        return dz * grad.to(dz.dtype), None, None, None


def repa_cosine_loss(z, y, grid_hw, teacher_hw=(16, 16)):
    """L = -mean_{b,t} cos(avgpool(z)[b,t], y[b,t]);  z [B, gh*gw, E] bf16, y [B, th*tw, E]."""
    return _RepaCos.apply(z, y, tuple(grid_hw), tuple(teacher_hw))


class RepaModel(torch.nn.Module):
    """Base model + REPA projection heads (common/repa.py:82-135)."""

    def __init__(self, model: torch.nn.Module, repa_config: RepaConfig) -> None:
        super().__init__()
        self.model = model
        self.config = getattr(model, "config", None)
        self.repa_states = []
        self.repa_config = repa_config
        self.parent_modules = {}
        self._patch_base_model()

    def _patch_base_model(self):
        out_size = self.repa_config.target_shape
        for idx, module_name in enumerate(self.repa_config.target_modules):
            hidden_shape = self.repa_config.hidden_shape[idx]
            module = self.model.get_submodule(module_name)
            ref = next(module.parameters(), None)
            patched = RepaModule(module, hidden_shape, out_size)
            if ref is not None:
                patched.proj_mlp.to(device=ref.device, dtype=ref.dtype)
            self.model.set_submodule(module_name, patched)

    def get_base_model(self):
        for module_name in self.repa_config.target_modules:
            module = self.model.get_submodule(module_name)
            self.model.set_submodule(module_name, module.base_module)
        return self.model

    def forward(self, *args, **kwargs):
        return self.model(*args, **kwargs)

    def calculate_loss(self, features, grid_hw, teacher_hw=(16, 16)):
        """Mean over target modules of the cosine alignment loss; resets the stashed states."""
        total = None
        for module_name in self.repa_config.target_modules:
            module = self.model.get_submodule(module_name)
            zs = module.zs
            loss = repa_cosine_loss(zs.to(torch.bfloat16), features, grid_hw, teacher_hw)
            total = loss if total is None else total + loss
            module.zs = None
        return total / len(self.repa_config.target_modules)

    def save_pretrained(self, save_directory):
        """Base model via its own ``save_pretrained`` + ``repa_config.json`` +
        ``repa_model_weights.bin`` (common/repa.py:156-186).  The trained projector modules are
        re-attached as they are (the reference re-creates them here, losing their weights)."""
        os.makedirs(save_directory, exist_ok=True)
        patched = {name: self.model.get_submodule(name) for name in self.repa_config.target_modules}
        heads = {name: m.proj_mlp.state_dict() for name, m in patched.items()}
        original_model = self.get_base_model()
        try:
            if hasattr(original_model, "save_pretrained"):
                original_model.save_pretrained(save_directory)
        finally:
            for name, m in patched.items():
                self.model.set_submodule(name, m)
        with open(os.path.join(save_directory, "repa_config.json"), "w") as f:
            json.dump(self.repa_config.__dict__, f, indent=4)
        torch.save(heads, os.path.join(save_directory, "repa_model_weights.bin"))

    @classmethod
    def from_pretrained(cls, base_model, pretrained_model_name_or_path, **kwargs):
        config_path = os.path.join(pretrained_model_name_or_path, "repa_config.json")
        if not os.path.exists(config_path):
            raise FileNotFoundError(f"Repa config file not found at {config_path}")
        with open(config_path, "r") as f:
            repa_config = RepaConfig(**json.load(f))
        repa_model = cls(base_model, repa_config)
        weights = os.path.join(pretrained_model_name_or_path, "repa_model_weights.bin")
        if os.path.exists(weights):
            sd = torch.load(weights, map_location="cpu")
            for module_name, state_dict in sd.items():
                repa_model.model.get_submodule(module_name).proj_mlp.load_state_dict(state_dict)
        return repa_model
