"""A/B of step-level switches in ONE process on one box (box-to-box clock variance under the power cap is 1-2 %):
SANA-1.6B batch 16, resident synthetic batches, CUDA events, configurations interleaved."""
import sys
import torch
sys.path.insert(0, ".")
from train_sana import SanaModel
from yat_b200.attention import AttnProcessor2_0, SanaLinearAttnProcessorB200
from yat_b200.training_parameters_reader import TrainingParameters
p = TrainingParameters()
p.batch_size = 16
p.learning_rate = 1e-5
p.model_config = "sana_1600m_1024"
tr = SanaModel(p)
tr.initialize()
it = iter(tr.sampler)
batches = [next(it) for _ in range(2)]
procs = [b.attn1.processor for b in tr.pipe.transformer.transformer_blocks]


def setcfg(fuse, fold):
    for pr in procs:
        if isinstance(pr, SanaLinearAttnProcessorB200):
            pr.fuse_qkv = fuse
    AttnProcessor2_0.fold_key_mask = fold


def run(n):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        tr.train_step(batches[i % 2])
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


cfgs = {"fuse+fold": (True, True), "nofuse+fold": (False, True), "fuse+nofold": (True, False), "nofuse+nofold": (False, False)}
for name, c in cfgs.items():
    setcfg(*c)
    run(3)
res = {k: [] for k in cfgs}
for rep in range(3):
    for name, c in cfgs.items():
        setcfg(*c)
        run(1)
        res[name].append(run(5))
for k, v in res.items():
    print(f"{k:14s} " + " ".join(f"{x:7.2f}" for x in v) + f"   min {min(v):.2f} ms")
