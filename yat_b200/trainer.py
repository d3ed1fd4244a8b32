"""Trainer core: the ``Model`` hook surface and the step loop of common/trainer.py.

Same hook names and signatures as the reference (``extract_latents``, ``extract_embeddings``,
``format_embeddings``, ``get_timesteps``, ``find_closest_ratio``, ``enable_efficient_attention``,
``initialize``, ``validate``, ``optimize(ratio, latents, embeddings, repa_tokens, generator)``,
``repa_loss``, ``save_model``, ``finalize``, ``run``) and the same attributes the loop reads
(``model pipe scheduler aspect_ratios accelerator params optimizer sampler empty_embeddings
global_step``), common/trainer.py:93-124,283-403.  Carried over from the step loop: the PEFT wrap for LoRA / LoKr
(:212-241, ``adapters.py``), AdamW + global-norm clip (:245-248,347-348, ``optim.py``), LR warm-up (:255-262), EMA
(:264-268,350-351,374-377,381-400, ``ema.py``), CFG dropout with the empty-caption embedding (:308,319-323),
exploration (:326-336), the loss gather (:359).  What is *not* carried over: TensorBoard / validation image sampling /
shard download (out of the hot-path scope), the NCCL_P2P_DISABLE env (common/trainer.py:27-29), and
``raise NotImplemented`` (a TypeError upstream; here NotImplementedError).  Config keys this build parses but does
not implement (LoHa / FourierFT / DoRA / lora_dropout / lora_pretrained / use_adamw_8bit) raise at ``initialize``
instead of training a different model silently.
"""
import os
import random

import torch
from torch.optim.lr_scheduler import LambdaLR

from .accelerator import Accelerator
from .aspect_ratios import ASPECT_RATIO_512_BIN, ASPECT_RATIO_1024_BIN
from .bucket_sampler import SyntheticBucketSampler, find_closest_ratio
from .noise import PhiloxGenerator
from .training_parameters_reader import TrainingParameters


class Model:
    def __init__(self, params: TrainingParameters):
        self.accelerator = Accelerator(gradient_accumulation_steps=params.gradient_accumulation_steps)
        self.params = params
        self.process_index = self.accelerator.process_index
        self.num_processes = self.accelerator.num_processes
        self.timesteps = [int(t) for t in params.timesteps]
        if len(self.timesteps) != 0:
            # refiner-style training on a fixed list of timestep indices (common/trainer.py:51-64)
            def get_timesteps_from_list(batch_size):
                indices = [random.choice(self.timesteps) for _ in range(batch_size)]
                return self.scheduler.timesteps[indices].to(self.accelerator.device)
            self.get_timesteps = get_timesteps_from_list
        # rank-disjoint shard ranges (common/trainer.py:66-84)
        n = params.num_shards or 0
        if n >= self.num_processes and params.dreambooth_dataset_folder is None:
            per = n // self.num_processes
            self.shard_index_begin = self.process_index * per
            self.shard_index_end = self.shard_index_begin + per if self.process_index != self.num_processes - 1 else n
        else:
            self.shard_index_begin, self.shard_index_end = 0, n
        self.global_step = 0
        self.model = None
        self.pipe = None
        self.scheduler = None
        self.aspect_ratios = ASPECT_RATIO_1024_BIN
        self.optimizer = None
        self.sampler = None
        self.lr_scheduler = None
        self.ema_model = None
        self.logger = None
        self.empty_embeddings = None
        self.losses = []

    # ---- hooks -------------------------------------------------------------------------
    def extract_latents(self, images):
        raise NotImplementedError

    def extract_embeddings(self, captions):
        raise NotImplementedError

    def format_embeddings(self, embeds):
        pass

    def get_timesteps(self, batch_size):
        """Logit-normal timestep draw (common/trainer.py:96-101) through the fused kernel's
        sampler: returns scheduler timesteps for ``batch_size`` fresh indices."""
        from .noise import flowmatch_noise
        dummy = torch.zeros((batch_size, 8), dtype=torch.bfloat16, device=self.accelerator.device)
        seed, off = self._generator().advance()
        return flowmatch_noise(dummy, self.scheduler, seed=seed, offset=off, want_target=False).timesteps

    def find_closest_ratio(self, ratio):
        return find_closest_ratio(self.aspect_ratios, ratio)

    def enable_efficient_attention(self):
        pass  # SDPA is already the cross-attention path; linear attention is the tcgen05 kernel

    def validate(self):
        raise NotImplementedError

    def optimize(self, ratio, latents, embeddings, repa_tokens, generator):
        raise NotImplementedError

    def finalize(self):
        pass

    def repa_loss(self, batch):
        return 0.0

    def save_model(self):
        """common/trainer.py:295-296.  With adapters attached the reference saves a PeftModel, i.e. the adapter
        tensors + their config only; the frozen base weights are not rewritten."""
        model = self.accelerator.unwrap_model(self.model)
        path = f"models/{self.global_step}"
        if self.params.lora_rank is not None:
            from .adapters import save_adapters
            save_adapters(model, path, algo=self.params.lora_algo or "lora", rank=self.params.lora_rank,
                          alpha=self.params.lora_alpha, target_modules=self.params.lora_target_modules)
        else:
            model.save_pretrained(path)

    def make_sampler(self):
        """Sampler of ``Batch`` records.  Synthetic by default (dataset plumbing is out of scope)."""
        raise NotImplementedError

    # ---- plumbing ------------------------------------------------------------------------
    def _generator(self):
        if not hasattr(self, "_philox"):
            self._philox = PhiloxGenerator(rank=self.process_index)
        return self._philox

    def initialize(self):
        params = self.params
        if params.aspect_ratios == 512:
            self.aspect_ratios = ASPECT_RATIO_512_BIN
        elif params.aspect_ratios == 1024:
            self.aspect_ratios = ASPECT_RATIO_1024_BIN
        self.enable_efficient_attention()
        if self.accelerator.device.type != "cuda":
            raise RuntimeError("yat_b200.trainer.Model.initialize needs a CUDA device: the optimiser, EMA and every "
                               "hot-path op are sm_100a kernels (no CPU fallback)")
        if self.accelerator.is_main_process:
            os.makedirs("models", exist_ok=True)
        self.sampler = self.make_sampler()
        if params.use_adamw_8bit:
            raise NotImplementedError("use_adamw_8bit (bitsandbytes Lion8bit, common/trainer.py:249-250) is not built")
        if params.lora_rank is not None:                                   # PEFT wrap (common/trainer.py:212-241)
            from .adapters import apply_adapters
            if params.lora_pretrained is not None:
                raise NotImplementedError("lora_pretrained: loading a saved PEFT adapter is outside the hot-path scope")
            if params.lora_dropout:
                raise NotImplementedError(f"lora_dropout={params.lora_dropout}: adapter dropout is not built; "
                                          "set it to 0 (common/trainer.py:216,221,226)")
            if params.lora_use_dora:
                raise NotImplementedError("lora_use_dora: DoRA is not built (common/trainer.py:219)")
            base = self.pipe.transformer if self.pipe is not None and getattr(self.pipe, "transformer", None) is not None \
                else self.model
            self.adapted_modules = apply_adapters(base, params.lora_algo or "lora", params.lora_rank, params.lora_alpha,
                                                  params.lora_target_modules)
        trainable = [p for p in self.model.parameters() if p.requires_grad]
        # clip_grad_norm_ + AdamW.step as two persistent kernels (common/trainer.py:245-248,347-348)
        from .optim import FusedClipAdamW
        self.optimizer = FusedClipAdamW(trainable, lr=params.learning_rate, weight_decay=params.weight_decay)
        self.model = self.accelerator.prepare(self.model)
        if params.warmup_steps is not None:
            warm = params.warmup_steps
            self.lr_scheduler = LambdaLR(self.optimizer, lambda s: float(s) / float(max(1, warm)) if s < warm else 1.0)
        if params.use_ema:                                                 # common/trainer.py:264-268
            from .ema import EMAModel
            plist = list(self.model.parameters())
            self.ema_model = EMAModel(plist, decay=0.999).bind(plist)
        if params.train_unconditional_prob > 0 and self.empty_embeddings is None:
            # common/trainer.py:308 encodes [''] before the loop; the text encoder is out of scope here, so the hook
            # (or the latent cache) has to supply the empty-caption embedding
            self.empty_embeddings = self.make_empty_embeddings()
            if self.empty_embeddings is None:
                raise ValueError("train_unconditional_prob > 0 needs the empty-caption embedding: set "
                                 "trainer.empty_embeddings (list with one [L, E] tensor) or override "
                                 "make_empty_embeddings(); the reference computes extract_embeddings(['']) "
                                 "(common/trainer.py:308)")

    def make_empty_embeddings(self):
        """Embedding of the empty caption for CFG dropout (common/trainer.py:308), or None if unavailable."""
        try:
            return self.extract_embeddings([""])
        except NotImplementedError:
            return None

    def train_step(self, batch):
        """One micro-step of common/trainer.py:313-356; returns the (detached) loss tensor."""
        params = self.params
        ratio, latents, embeddings, repa_features = batch.ratio, batch.vae_features, batch.embeddings, batch.repa_features
        with self.accelerator.accumulate(self.model):
            if random.random() < params.train_unconditional_prob and self.empty_embeddings is not None:
                embeddings = [self.empty_embeddings[0] for _ in embeddings]       # CFG dropout (:319-323)
            generator = self._generator()                                          # (:325) caller-owned RNG state
            if params.exploration_steps is not None:                               # (:326-336)
                with torch.no_grad():
                    states, losses = [], []
                    for _ in range(params.exploration_steps):
                        states.append(generator.get_state())
                        losses.append(self.optimize(ratio, latents, embeddings, repa_features, generator))
                    generator.set_state(states[int(torch.argmin(torch.stack(losses)))])
            loss = self.optimize(ratio, latents, embeddings, repa_features, generator)
            self.accelerator.backward(loss)
            if self.accelerator.sync_gradients:
                # (:347-351) clip_grad_norm_(1.0); optimizer.step(); ema_model.step(parameters): two launches
                self.optimizer.clip_and_step(max_norm=1.0, ema=self.ema_model)
                if self.lr_scheduler is not None:
                    self.lr_scheduler.step()
                self.optimizer.zero_grad(set_to_none=True)
        return loss.detach()

    def run(self):
        params = self.params
        self.initialize()
        self.accelerator.wait_for_everyone()
        avg_loss = torch.zeros((), device=self.accelerator.device)
        while self.global_step < params.steps:
            for batch in self.sampler:
                avg_loss = avg_loss + self.train_step(batch).float()
                if self.accelerator.sync_gradients:
                    mean_loss = self.accelerator.gather(avg_loss).mean()          # (:359)
                    avg_loss = torch.zeros((), device=self.accelerator.device)
                    self.losses.append(mean_loss)
                    if self.global_step % params.num_steps_per_validation == 0 and self.global_step > 0:
                        with torch.no_grad():
                            if self.ema_model is not None:                        # (:374-377) on ALL ranks
                                self.ema_model.reduce_mean(self.accelerator)
                            if self.accelerator.is_main_process:
                                if self.ema_model is not None:                    # (:381-383)
                                    self.ema_model.store(self.model.parameters())
                                    self.ema_model.copy_to(self.model.parameters())
                                self.save_model()                                 # (:398; validate() is out of scope)
                                if self.ema_model is not None:                    # (:399-400)
                                    self.ema_model.restore(self.model.parameters())
                    self.global_step += 1
                    if self.global_step >= params.steps:
                        break
        self.finalize()
