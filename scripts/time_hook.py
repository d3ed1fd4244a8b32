"""Train-step time of a hook at its BASELINE.json config (K3 PixArt-Sigma XL / K4 SD3.5-medium), resident batches.
Not the headline bench (bench.py is K2); a parity-config sanity number for DESIGN.md."""
import sys
import time
import torch
sys.path.insert(0, ".")
from yat_b200.training_parameters_reader import TrainingParameters
which, batch, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
p = TrainingParameters()
p.batch_size = batch
p.learning_rate = 1e-5
if which == "pixart":
    from train_pixart_sigma import PixartSigmaTrainer as T
    p.model_config = "pixart_sigma_xl_1024"
elif which == "sana_k5":                  # BASELINE config 5: SANA-1.6B + REPA + use_preservation, LoKr rank 16
    from train_sana import SanaModel as T
    p.model_config = "sana_1600m_1024"
    p.use_repa, p.use_preservation, p.preservation_ratio = True, True, 0.5
    p.lora_rank, p.lora_alpha, p.lora_algo = 16, 16, "lokr"
    p.lora_target_modules = "to_q to_k to_v to_out.0 conv_inverted conv_point linear_1 linear_2 proj"   # README.md:78
    p.learning_rate = 1e-4
else:
    from train_sd35 import SD35Trainer as T
    p.model_config = "sd35_medium"
    if which == "sd35_lora":              # BASELINE config 4: LoRA rank 32
        p.lora_rank, p.lora_alpha, p.lora_algo = 32, 32, "lora"
        p.lora_target_modules = "to_q to_k to_v to_out.0 add_q_proj add_k_proj add_v_proj to_add_out"
        p.learning_rate = 1e-4
tr = T(p)
tr.initialize()
it = iter(tr.sampler)
batches = [next(it) for _ in range(2)]
for i in range(3):
    loss = tr.train_step(batches[i % 2])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(steps):
    loss = tr.train_step(batches[i % 2])
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
print(f"{which} batch {batch}: {ms:.1f} ms/step, {batch / ms * 1e3:.1f} img/s, loss {loss.item():.4f}, "
      f"mem {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
