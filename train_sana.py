"""SANA trainer hooks on the B200 hot path — same class and hook names as the reference's
train_sana.py (``SanaModel(Model)``: ``extract_latents`` / ``extract_embeddings`` /
``enable_efficient_attention`` / ``validate`` / ``optimize``), so a config that ran there runs here.

``optimize`` is the flow-matching step of train_sana.py:163-219 rebuilt on three fused ops:
  pad/mask of the ragged text embeddings on the device      (:167-179  -> yat_pad_embeddings)
  noise + logit-normal timestep + sigma lookup + x_t         (:183-207  -> yat_noise_flowmatch)
  target + fp32 MSE + its backward                           (:217-218  -> yat_loss_fused)
with SANA's linear attention inside the model running as tcgen05 kernels.  No ``.item()`` and no
host sync anywhere in the step.

The VAE / text-encoder / pipeline halves of the reference class (``SanaPipeline.from_pretrained``,
``validate`` image sampling) need diffusers + weights and are out of this build's scope; the
transformer is built from ``params.model_config`` (random init, or ``pretrained_model_path`` with a
HF-layout ``config.json`` + safetensors).
"""
import argparse
import copy
from types import SimpleNamespace

import torch

from yat_b200.aspect_ratios import (ASPECT_RATIO_256_BIN, ASPECT_RATIO_512_BIN, ASPECT_RATIO_1024_BIN,
                                    ASPECT_RATIO_2048_BIN)
from yat_b200.bucket_sampler import SyntheticBucketSampler
from yat_b200.embeddings import pad_embeddings
from yat_b200.loss import VELOCITY, fused_denoise_loss, sd3_loss_weights
from yat_b200.noise import flowmatch_noise
from yat_b200.repa import RepaConfig, RepaModel
from yat_b200.sana_transformer import SanaTransformer2DModel
from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
from yat_b200.trainer import Model
from yat_b200.training_parameters_reader import TrainingParameters

SANA_CONFIGS = {
    # utils/patched_sana_transformer.py:88-112 defaults = SANA-1.6B 1024px
    "sana_1600m_1024": dict(),
    "sana_600m_1024": dict(num_attention_heads=36, num_layers=28, num_cross_attention_heads=16,
                           cross_attention_head_dim=72, cross_attention_dim=1152),
    # BASELINE.json config 1: tiny random-init 2-layer model, 32-ch 16x16 latents
    "sana_tiny": dict(num_attention_heads=4, attention_head_dim=32, num_layers=2, num_cross_attention_heads=2,
                      cross_attention_head_dim=64, cross_attention_dim=128, caption_channels=64, sample_size=16),
}


class SanaModel(Model):
    max_text_tokens = 512
    # pad the text batch only up to its longest caption (rounded up to 64 tokens) instead of always 512:
    # identical cross-attention result (masked keys weigh exactly 0), ~1.6x less K/V projection + SDPA work
    trim_text_padding = 64

    def __init__(self, params: TrainingParameters):
        super().__init__(params)
        device = self.accelerator.device
        if params.pretrained_model_path is not None:
            transformer = SanaTransformer2DModel.from_pretrained(params.pretrained_model_path,
                                                                 torch_dtype=torch.bfloat16, device_map=device)
        else:
            cfg = params.model_config or "sana_1600m_1024"
            cfg = SANA_CONFIGS[cfg] if isinstance(cfg, str) else cfg
            with torch.device(device):
                transformer = SanaTransformer2DModel(**cfg).to(torch.bfloat16)
        self.pipe = SimpleNamespace(transformer=transformer, vae=None, text_encoder=None)
        shift = float(getattr(params, "flow_shift", 1.0) or 1.0)
        self.scheduler = FlowMatchEulerDiscreteScheduler(shift=shift)
        resolution = transformer.config.sample_size * 32                     # train_sana.py:45-54
        self.aspect_ratios = {256: ASPECT_RATIO_256_BIN, 512: ASPECT_RATIO_512_BIN,
                              1024: ASPECT_RATIO_1024_BIN}.get(resolution, ASPECT_RATIO_2048_BIN)
        self.model = transformer
        # train_sana.py:63 turns gradient checkpointing on unconditionally (a 24-48 GB-card habit); with
        # 180 GB of HBM3e the saved activations of SANA-1.6B at batch 16 (~50 GB) fit, so "auto" keeps them
        # and skips the recompute forward.  Numerically identical either way.
        if self._want_checkpointing(transformer, params):
            self.model.enable_gradient_checkpointing()
        self.frozen_model = None
        self.patch_size = transformer.config.patch_size
        self._trust_gscale = False
        if params.use_repa:
            blk = f"transformer_blocks.{min(7, transformer.config.num_layers - 1)}"
            inner = transformer.config.num_attention_heads * transformer.config.attention_head_dim
            self.model = RepaModel(transformer, RepaConfig(target_modules=[blk], hidden_shape=[inner]))
        if params.use_preservation and params.lora_rank is None:
            self.frozen_model = copy.deepcopy(transformer).requires_grad_(False).eval()
        # with LoRA / LoKr the frozen prediction is the same network with its adapters disabled

    def _want_checkpointing(self, transformer, params):
        mode = getattr(params, "gradient_checkpointing", "auto")
        if mode != "auto":
            return bool(mode)
        cfg = transformer.config
        if self.accelerator.device.type != "cuda":
            return True
        tokens = (cfg.sample_size // cfg.patch_size) ** 2
        inner = cfg.num_attention_heads * cfg.attention_head_dim
        # ~35 [B, N, C] bf16 tensors are kept per block (projections, attention, the 5x-wide GLUMBConv)
        need = 35 * params.batch_size * tokens * inner * 2 * cfg.num_layers * 1.25
        n_params = sum(p.numel() for p in transformer.parameters())
        need += n_params * 2 * 5 + (8 << 30)                 # weights, grads, 2 Adam moments, DDP buckets, workspace
        free, _ = torch.cuda.mem_get_info(self.accelerator.device)
        return need > free

    def initialize(self):
        super().initialize()
        self.pipe.transformer = self.accelerator.unwrap_model(self.model)

    def make_sampler(self):
        cfg = self.pipe.transformer.config
        s = cfg.sample_size
        if getattr(self.params, "latent_cache", None):
            # per-bucket bf16 slabs: every rank reads the same index, so all ranks step on the
            # same bucket without the reference's barrier + all_gather per element (common/bucket_sampler.py:225-237)
            from yat_b200.latent_cache import SlabBucketSampler, SlabReader
            return SlabBucketSampler(SlabReader(self.params.latent_cache), self.params.batch_size,
                                     rank=self.process_index, world=self.num_processes, seed=1234)
        return SyntheticBucketSampler(self.params.batch_size, (cfg.in_channels, s, s), cfg.caption_channels,
                                      ratio=1.0, seed=1234 + self.process_index, repa_tokens=self.params.use_repa)

    def extract_latents(self, images):
        raise NotImplementedError("DC-AE encoding needs the SANA pipeline weights (out of scope); "
                                  "feed pre-computed latents (train_sana.py:78-82 multiplies by scaling_factor)")

    def extract_embeddings(self, captions):
        raise NotImplementedError("Gemma prompt encoding needs the SANA pipeline weights (out of scope)")

    def enable_efficient_attention(self):
        pass

    def train_step(self, batch):
        self._trust_gscale = True   # the trainer owns backward(): upstream grad is exactly 1/accum
        try:
            return super().train_step(batch)
        finally:
            self._trust_gscale = False

    def pad_text(self, embeddings):
        """Ragged caption embeddings -> (padded [B, L, E] bf16, mask int64 [B, L]) on the device."""
        return pad_embeddings(embeddings, self.max_text_tokens, self.accelerator.device,
                              trim_multiple=self.trim_text_padding)

    def optimize(self, ratio, latents, embeddings, repa_tokens, generator=None):
        params = self.params
        device = self.accelerator.device

        # (1) text: pad to 512 tokens + mask, on the device
        if isinstance(embeddings, tuple):
            embeddings, attention_mask = embeddings
        else:
            embeddings, attention_mask = self.pad_text(embeddings)
        latents = latents.to(device=device, dtype=torch.bfloat16, non_blocking=True)

        # (2) noising.  A torch.Generator reproduces the reference draw for draw (noise, then u,
        # from the same CPU generator, train_sana.py:183-191) and injects them; otherwise eps and u
        # come from the counter RNG inside the kernel and eps never touches HBM.
        if isinstance(generator, torch.Generator):
            noise = torch.randn(latents.shape, generator=generator, device="cpu", dtype=torch.bfloat16)
            u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(latents.shape[0],), generator=generator))
            nz = flowmatch_noise(latents, self.scheduler, noise=noise.to(device), u=u.to(device), want_target=False)
            seed = offset = 0
        else:
            seed, offset = (generator or self._generator()).advance()
            nz = flowmatch_noise(latents, self.scheduler, seed=seed, offset=offset, want_target=False)

        # (3) model (bf16 end to end)
        noise_pred = self.model(nz.x_t, encoder_hidden_states=embeddings, timestep=nz.timesteps,
                                encoder_attention_mask=attention_mask).sample

        # (4) loss: fp32 velocity MSE + optional preservation, forward and dL/dpred in one kernel
        pred_frozen = None
        if self.frozen_model is not None:
            with torch.no_grad():
                pred_frozen = self.frozen_model(nz.x_t, encoder_hidden_states=embeddings, timestep=nz.timesteps,
                                                encoder_attention_mask=attention_mask).sample
        elif params.use_preservation and params.lora_rank is not None:
            from yat_b200.adapters import adapters_disabled
            with torch.no_grad(), adapters_disabled(self.pipe.transformer):
                pred_frozen = self.pipe.transformer(nz.x_t, encoder_hidden_states=embeddings, timestep=nz.timesteps,
                                                    encoder_attention_mask=attention_mask).sample
        loss = fused_denoise_loss(noise_pred, latents, nz.noise, pred_frozen=pred_frozen,
                                  weights=sd3_loss_weights(getattr(params, "weighting_scheme", "none"), nz.sigmas),
                                  rho=float(params.preservation_ratio or 0.0),
                                  gscale=1.0 / self.accelerator.gradient_accumulation_steps,
                                  target_mode=VELOCITY, lowp=False, seed=seed, offset=offset,
                                  trust_gscale=self._trust_gscale)
        if params.use_repa and repa_tokens is not None:
            h, w = latents.shape[-2] // self.patch_size, latents.shape[-1] // self.patch_size
            repa = self.accelerator.unwrap_model(self.model).calculate_loss(
                repa_tokens.to(device, non_blocking=True), (h, w))
            loss = loss + params.repa_lambda * repa
        return loss


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--config", required=True, type=str)
    args = parser.parse_args()
    params = TrainingParameters()
    params.read_yaml(args.config)
    trainer = SanaModel(params)
    trainer.run()
