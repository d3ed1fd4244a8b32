"""SD3.5 trainer hooks on the B200 hot path — same class and hook names as the reference's train_sd35.py
(``SD35Trainer(Model)``: ``extract_latents`` / ``extract_embeddings`` / ``validate`` / ``optimize``).

``optimize`` is the flow-matching step of train_sd35.py:165-194 rebuilt on the fused ops:
  noise + logit-normal index + sigma lookup (shift 3.0) + scale_noise   (:180-185 -> yat_noise_flowmatch)
  velocity target noise - latents, bf16 MSE                              (:192-193 -> yat_loss_fused, lowp)
A batch is ``(latents, embeddings, pooled_projections)`` (:177).  The SD3.5-medium MMDiT is
``yat_b200.sd3_transformer.SD3Transformer2DModel`` (the diffusers architecture restated with the HF
key layout on the fused block ops; random-init of ``params.model_config`` when no module is passed); any module with
the same forward signature can be passed as ``transformer`` (``yat_b200.toy_models.ToySD3Transformer`` is a small
stand-in some tests use).  LoRA (rank 32 in BASELINE.json config 4) is the adapter wrap of the base trainer
(``yat_b200.adapters``, ``lora_rank`` / ``lora_target_modules``) around that transformer and does not touch this path.
"""
import argparse
from types import SimpleNamespace

import torch

from yat_b200.aspect_ratios import ASPECT_RATIO_1024_BIN
from yat_b200.bucket_sampler import Batch, SyntheticBucketSampler
from yat_b200.loss import VELOCITY, fused_denoise_loss, sd3_loss_weights
from yat_b200.noise import flowmatch_noise
from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
from yat_b200.sd3_transformer import SD3_CONFIGS, SD3Transformer2DModel
from yat_b200.trainer import Model
from yat_b200.training_parameters_reader import TrainingParameters


class _SD35Sampler(SyntheticBucketSampler):
    """Synthetic SD3.5 batches: fixed-length joint text embeddings [B, 333, 4096] + pooled [B, 2048]."""

    def __init__(self, batch_size, latent_shape, caption_channels, pooled_channels, seed):
        super().__init__(batch_size, latent_shape, caption_channels, ratio=1.0, seed=seed, max_len=333)
        g = torch.Generator().manual_seed(seed + 1)
        for b in self.batches:
            emb = (0.5 * torch.randn((batch_size, 333, caption_channels), generator=g)).to(torch.bfloat16)
            pooled = (0.5 * torch.randn((batch_size, pooled_channels), generator=g)).to(torch.bfloat16)
            b.embeddings = (emb, pooled)


class SD35Trainer(Model):
    latent_channels = 16
    vae_scale = 8

    def __init__(self, params: TrainingParameters, transformer=None):
        super().__init__(params)
        if transformer is None:
            # random-init SD3.5 MMDiT of the named size (BASELINE.json config 4; there are no weights offline).  A
            # checkpoint of the reference's pipeline loads into it with load_state_dict (HF key layout).
            cfg = params.model_config or "sd35_medium"
            cfg = SD3_CONFIGS[cfg] if isinstance(cfg, str) else cfg
            with torch.device(self.accelerator.device):
                transformer = SD3Transformer2DModel(**cfg)
        self.pipe = SimpleNamespace(transformer=transformer, vae=None, text_encoder=None, text_encoder_2=None,
                                    text_encoder_3=None)
        shift = float(getattr(params, "flow_shift", 3.0) or 3.0)          # SD3.5 scheduler config: shift 3.0
        self.scheduler = FlowMatchEulerDiscreteScheduler(shift=shift)
        self.aspect_ratios = ASPECT_RATIO_1024_BIN
        self.model = transformer.to(self.accelerator.device, torch.bfloat16)
        if hasattr(self.model, "enable_gradient_checkpointing") and params.gradient_checkpointing is True:
            self.model.enable_gradient_checkpointing()
        self._trust_gscale = False

    def initialize(self):
        super().initialize()
        self.pipe.transformer = self.accelerator.unwrap_model(self.model)

    def make_sampler(self):
        cfg = getattr(self.pipe.transformer, "config", SimpleNamespace())
        side = getattr(cfg, "sample_size", None) or 1024 // self.vae_scale
        return _SD35Sampler(self.params.batch_size, (self.latent_channels, side, side),
                            getattr(cfg, "caption_channels", 4096), 2048, 1234 + self.process_index)

    def extract_latents(self, images):
        raise NotImplementedError("SD3 VAE encoding needs the pipeline weights (out of scope); feed cached latents")

    def extract_embeddings(self, caption):
        raise NotImplementedError("CLIP-L/CLIP-G/T5 prompt encoding needs the pipeline weights (out of scope)")

    def train_step(self, batch):
        self._trust_gscale = True
        try:
            return super().train_step(batch)
        finally:
            self._trust_gscale = False

    def optimize(self, ratio, latents=None, embeddings=None, repa_tokens=None, generator=None):
        """Accepts the base trainer's call (ratio, latents, (embeddings, pooled), repa_tokens, generator) and
        the reference hook's own (model, batch) form with batch = (latents, embeddings, pooled)."""
        if isinstance(latents, (tuple, list)) and len(latents) == 3:           # optimize(model, batch)
            latents, embeddings, pooled = latents
        else:
            embeddings, pooled = embeddings
        device = self.accelerator.device
        latents = latents.to(device=device, dtype=torch.bfloat16, non_blocking=True)
        embeddings = embeddings.to(device=device, dtype=torch.bfloat16, non_blocking=True)
        pooled = pooled.to(device=device, dtype=torch.bfloat16, non_blocking=True)
        if isinstance(generator, torch.Generator):
            noise = torch.randn(latents.shape, generator=generator, device="cpu", dtype=torch.bfloat16)
            u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(latents.shape[0],), generator=generator))
            nz = flowmatch_noise(latents, self.scheduler, noise=noise.to(device), u=u.to(device), want_target=False)
            seed = offset = 0
        else:
            seed, offset = (generator or self._generator()).advance()
            nz = flowmatch_noise(latents, self.scheduler, seed=seed, offset=offset, want_target=False)
        pred = self.model(nz.x_t, encoder_hidden_states=embeddings, timestep=nz.timesteps,
                          pooled_projections=pooled).sample
        return fused_denoise_loss(pred.to(torch.bfloat16), latents, nz.noise,
                                  weights=sd3_loss_weights(getattr(self.params, "weighting_scheme", "none"), nz.sigmas),
                                  gscale=1.0 / self.accelerator.gradient_accumulation_steps, target_mode=VELOCITY,
                                  lowp=True, seed=seed, offset=offset, trust_gscale=self._trust_gscale)


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--config", required=True, type=str)
    args = parser.parse_args()
    params = TrainingParameters()
    params.read_yaml(args.config)
    trainer = SD35Trainer(params)
    trainer.run()
