"""Kernel-time breakdown of one K3 step (PixArt-Sigma XL 1024 px, batch 16), torch.profiler."""
import sys
import torch
sys.path.insert(0, ".")
from torch.profiler import ProfilerActivity, profile
from train_pixart_sigma import PixartSigmaTrainer as SanaModel
from yat_b200.training_parameters_reader import TrainingParameters
p = TrainingParameters()
p.batch_size = 16
p.model_config = "pixart_sigma_xl_1024"
p.learning_rate = 1e-4
tr = SanaModel(p)
tr.initialize()
it = iter(tr.sampler)
b = next(it)
for _ in range(3):
    tr.train_step(b)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    tr.train_step(b)
    torch.cuda.synchronize()
tot = 0
rows = []
for ev in prof.key_averages():
    if ev.device_type == torch.autograd.DeviceType.CUDA and ev.device_time_total > 0:
        rows.append((ev.device_time_total, ev.count, ev.key))
        tot += ev.device_time_total
print("total kernel ms", tot / 1e3)
for t, n, k in sorted(rows, reverse=True)[:28]:
    print(f"{t / 1e3:8.2f} ms x{n:5d}  {k[:120]}")
