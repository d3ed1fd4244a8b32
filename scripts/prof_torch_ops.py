"""Which torch ops launch the elementwise fill / add / copy kernels of a SANA step (tiny config, torch.profiler)."""
import collections
import sys
import torch
sys.path.insert(0, ".")
from torch.profiler import ProfilerActivity, profile
from train_sana import SanaModel
from yat_b200.training_parameters_reader import TrainingParameters
p = TrainingParameters()
p.batch_size = 2
p.model_config = "sana_tiny"
p.learning_rate = 1e-4
tr = SanaModel(p)
tr.initialize()
it = iter(tr.sampler)
b = next(it)
for _ in range(3):
    tr.train_step(b)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
    tr.train_step(b)
    torch.cuda.synchronize()
ev = prof.events()
# map each CUDA kernel to its launching CPU op (by correlation through time containment on the CPU side)
cpu_ops = [e for e in ev if e.device_type == torch.autograd.DeviceType.CPU]
cnt = collections.Counter()
for e in cpu_ops:
    ks = [k.name for k in e.kernels]
    for k in ks:
        if any(s in k for s in ("FillFunctor", "CUDAFunctor_add", "copy_kernel", "MulFunctor", "reduce_kernel")):
            shapes = str(e.input_shapes)[:70]
            par = e.cpu_parent.name if e.cpu_parent is not None else "-"
            gp = e.cpu_parent.cpu_parent.name if e.cpu_parent is not None and e.cpu_parent.cpu_parent is not None else "-"
            short = "Fill" if "Fill" in k else "add" if "_add" in k else "copy" if "copy" in k else "mul" if "Mul" in k else "reduce"
            cnt[(short, e.name, par[:40], gp[:40], shapes)] += 1
for key, n in sorted(cnt.items(), key=lambda t: (t[0][0], -t[1])):
    print(n, *key, sep=" | ")
