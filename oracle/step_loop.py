"""The step loop of the reference trainer on the CPU (oracle; test infrastructure only - nothing under yat_b200/ imports it).

Restates common/trainer.py:298-403 (`Model.run`) around train_sana.py:163-219 (`SanaModel.optimize`) for any callable
denoiser with the SanaTransformer2DModel signature:

  :311-317   for batch in sampler: with accelerator.accumulate(model)   micro-step counter, sync on every G-th
  :319-323   CFG dropout: ONE random.random() draw per micro-step; all captions of the batch -> the empty embedding
  :325       generator = torch.Generator()                              fresh and unseeded, every micro-step
  :326-336   exploration: E trial optimize() calls under no_grad, each from a saved generator state; the state of the
             lowest loss is restored, so the real step repeats that trial's draws
  :337       loss = optimize(...)                                       pad / mask, noise then u, x_t, model, fp32 MSE
  :343-344   avg_loss += loss (a SUM over the G micro-steps, not a mean); backward(loss / G)
  :346-356   on sync: clip_grad_norm_(1.0); AdamW.step(); lr_scheduler.step(); zero_grad()
  :358-367   logged train/loss = the summed loss, train/lr = get_last_lr()[0] after the scheduler step

Pinned by tests/test_reference_golden.py against tests/golden/reference_trainer_loop.pt, which was produced by executing
`Model.run` and `SanaModel.optimize` from the reference files themselves (tests/golden/make_golden_from_reference.py).
"""
import random

import torch
from torch.optim.lr_scheduler import LambdaLR

from . import losses, noising, sampling, schedule
from .train_step import pad_embeddings_reference


def optimize(model, latents, embeddings, generator, tables, batch_size):
    """train_sana.py:163-219 with the oracle's own pieces."""
    tt, ss = tables
    padded, mask = pad_embeddings_reference(embeddings, 512)
    latents = latents.to(torch.bfloat16)
    noise, u, idx = sampling.sana_step_draws(tuple(latents.shape), batch_size, torch.bfloat16, generator)
    x_t, target, t, _ = noising.flow_noise_sana(latents, noise, idx, tt, ss)
    pred = model(x_t, encoder_hidden_states=padded.to(torch.bfloat16), timestep=t, encoder_attention_mask=mask).sample
    return losses.mse_fp32(pred, target)


def run(model, batches, empty_embedding, steps, accum, batch_size, exploration_steps, uncond_prob, warmup_steps, lr,
        weight_decay, shift, random_seed):
    """Returns (logged losses, logged lrs); `model` is updated in place."""
    random.seed(random_seed)
    tables = schedule.flow_match_tables(1000, shift)
    opt = torch.optim.AdamW(model.parameters(), lr=lr, weight_decay=weight_decay)                       # :245-248
    sched = None
    if warmup_steps is not None:                                                                         # :255-262
        sched = LambdaLR(opt, lr_lambda=lambda s: float(s) / float(max(1, warmup_steps)) if s < warmup_steps else 1.0)
    logged, lrs, micro, global_step = [], [], 0, 0
    avg = torch.tensor(0)
    it = iter(batches)
    while global_step < steps:
        for batch in it:
            micro += 1
            sync = micro % accum == 0
            embeddings = batch.embeddings
            if random.random() < uncond_prob:
                embeddings = [empty_embedding for _ in embeddings]
            g = torch.Generator()
            if exploration_steps is not None:
                with torch.no_grad():
                    states, trial = [], []
                    for _ in range(exploration_steps):
                        states.append(g.get_state())
                        trial.append(optimize(model, batch.vae_features, embeddings, g, tables, batch_size))
                g.set_state(states[torch.argmin(torch.stack(trial))])
            loss = optimize(model, batch.vae_features, embeddings, g, tables, batch_size)
            avg = avg + loss
            (loss / accum).backward()
            if sync:
                torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
                opt.step()
                if sched is not None:
                    sched.step()
                opt.zero_grad()
                logged.append(float(avg.detach()))
                lrs.append(sched.get_last_lr()[0] if sched is not None else lr)
                avg = torch.tensor(0)
                global_step += 1
        break   # the golden's sampler holds exactly steps * accum micro-batches
    return logged, lrs
