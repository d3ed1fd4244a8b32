"""PixArt-Sigma trainer hooks on the B200 hot path — same class and hook names as the reference's
train_pixart_sigma.py (``PixartSigmaTrainer(Model)``: ``extract_latents`` / ``extract_embeddings`` /
``validate`` / ``optimize``).

``optimize`` is the DDPM eps-prediction step of train_pixart_sigma.py:151-185 rebuilt on the fused ops:
  pad/mask of the ragged T5 embeddings to 300 tokens              (:158-169 -> yat_pad_embeddings)
  noise + logit-normal index + alphas_cumprod lookup + add_noise   (:171-176 -> yat_noise_ddpm)
  eps target, first half of the 8 predicted channels, bf16 MSE     (:178-184 -> yat_loss_fused, EPS, lowp)
The PixArt-Sigma XL transformer is ``yat_b200.pixart_transformer.PixArtTransformer2DModel`` (the
diffusers architecture restated with the HF key layout on the fused block ops; random-init of ``params.model_config``
when no module is passed).  Any module with the same forward signature can be passed as ``transformer``
(``yat_b200.toy_models.ToyPixArtTransformer`` is a small stand-in some tests use).
"""
import argparse
from types import SimpleNamespace

import torch

from yat_b200.aspect_ratios import ASPECT_RATIO_1024_BIN
from yat_b200.bucket_sampler import SyntheticBucketSampler
from yat_b200.embeddings import pad_embeddings
from yat_b200.loss import EPS, fused_denoise_loss
from yat_b200.noise import ddpm_noise
from yat_b200.pixart_transformer import PIXART_CONFIGS, PixArtTransformer2DModel, REPAPixArtTransformerModel
from yat_b200.schedulers import DDPMScheduler
from yat_b200.trainer import Model
from yat_b200.training_parameters_reader import TrainingParameters


class PixartSigmaTrainer(Model):
    max_text_tokens = 300            # train_pixart_sigma.py:161
    latent_channels = 4
    vae_scale = 8                    # SDXL VAE: latent grid = pixels / 8

    def __init__(self, params: TrainingParameters, transformer=None):
        super().__init__(params)
        if transformer is None:
            # random-init PixArt-Sigma of the named size (BASELINE.json config 3; there are no weights offline).  A
            # checkpoint of the reference's pipeline loads into it with load_state_dict (HF key layout).
            cfg = params.model_config or "pixart_sigma_xl_1024"
            cfg = PIXART_CONFIGS[cfg] if isinstance(cfg, str) else cfg
            with torch.device(self.accelerator.device):
                # train_pixart_sigma.py:28-32: with use_repa the transformer is the REPA variant (projection head)
                transformer = REPAPixArtTransformerModel(**cfg) if params.use_repa else PixArtTransformer2DModel(**cfg)
        self.pipe = SimpleNamespace(transformer=transformer, vae=None, text_encoder=None, dtype=torch.bfloat16)
        self.scheduler = DDPMScheduler()                       # PixArt-Sigma: 1000 steps, linear betas 1e-4..0.02
        self.aspect_ratios = ASPECT_RATIO_1024_BIN
        self.model = transformer.to(self.accelerator.device, torch.bfloat16)
        if hasattr(self.model, "enable_gradient_checkpointing") and params.gradient_checkpointing is True:
            self.model.enable_gradient_checkpointing()         # train_pixart_sigma.py:55
        self._trust_gscale = False

    def initialize(self):
        super().initialize()
        self.pipe.transformer = self.accelerator.unwrap_model(self.model)

    def make_sampler(self):
        cfg = getattr(self.pipe.transformer, "config", SimpleNamespace())
        side = getattr(cfg, "sample_size", None) or 1024 // self.vae_scale
        return SyntheticBucketSampler(self.params.batch_size, (self.latent_channels, side, side),
                                      getattr(cfg, "caption_channels", 4096), ratio=1.0,
                                      seed=1234 + self.process_index, max_len=self.max_text_tokens,
                                      repa_tokens=bool(self.params.use_repa))

    def extract_latents(self, images):
        raise NotImplementedError("SDXL-VAE encoding needs the PixArt pipeline weights (out of scope); feed cached latents")

    def extract_embeddings(self, captions):
        raise NotImplementedError("T5 prompt encoding needs the PixArt pipeline weights (out of scope)")

    def train_step(self, batch):
        self._trust_gscale = True
        try:
            return super().train_step(batch)
        finally:
            self._trust_gscale = False

    def optimize(self, ratio, latents, embeddings=None, repa_tokens=None, generator=None):
        """Accepts the base trainer's call (ratio, latents, embeddings, repa_tokens, generator) and the
        reference hook's own (latents, embeddings) form."""
        if embeddings is None or not torch.is_tensor(latents):
            ratio, latents, embeddings = None, ratio, latents
        device = self.accelerator.device
        if isinstance(embeddings, tuple):
            embeddings, attention_mask = embeddings
        else:
            embeddings, attention_mask = pad_embeddings(embeddings, self.max_text_tokens, device)
        latents = latents.to(device=device, dtype=torch.bfloat16, non_blocking=True)
        if isinstance(generator, torch.Generator):
            # the reference draws noise with the device default generator and u on the CPU (:171-173)
            noise = torch.randn(latents.shape, generator=generator, device="cpu", dtype=torch.bfloat16)
            u = torch.sigmoid(torch.normal(mean=0.0, std=1.0, size=(latents.shape[0],), generator=generator))
            nz = ddpm_noise(latents, self.scheduler, noise=noise.to(device), u=u.to(device))
            seed = offset = 0
        else:
            seed, offset = (generator or self._generator()).advance()
            nz = ddpm_noise(latents, self.scheduler, seed=seed, offset=offset)
        pred = self.model(nz.x_t, encoder_hidden_states=embeddings, timestep=nz.timesteps,
                          encoder_attention_mask=attention_mask.to(torch.bfloat16)).sample
        # learned-sigma models predict 2C channels; the loss reads the first C (chunk(2, 1)[0], :182)
        loss = fused_denoise_loss(pred.to(torch.bfloat16), latents, nz.noise,
                                  gscale=1.0 / self.accelerator.gradient_accumulation_steps, target_mode=EPS,
                                  lowp=True, seed=seed, offset=offset, trust_gscale=self._trust_gscale)
        base = self.accelerator.unwrap_model(self.model)
        if self.params.use_repa and repa_tokens is not None and hasattr(base, "calculate_loss"):
            p = base.config.patch_size
            grid = (latents.shape[-2] // p, latents.shape[-1] // p)
            loss = loss + self.params.repa_lambda * base.calculate_loss(repa_tokens.to(device, non_blocking=True), grid)
        return loss


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("--config", required=True, type=str)
    args = parser.parse_args()
    params = TrainingParameters()
    params.read_yaml(args.config)
    trainer = PixartSigmaTrainer(params)
    trainer.run()
