"""One launch of every hot-path kernel at its roofline shape (for `ncu --set full`)."""
import sys
import torch
sys.path.insert(0, ".")
from yat_b200.linear_attention import sana_linear_attention
from yat_b200.loss import fused_denoise_loss
from yat_b200.noise import flowmatch_noise
from yat_b200.repa import repa_cosine_loss
from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
dev = torch.device("cuda")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
x0 = torch.randn((4096, 32, 32, 32), device=dev, dtype=torch.bfloat16)
pred = torch.randn_like(x0).requires_grad_(True)
B, N, H = 16, 1024, 70
q, k, v, do = [torch.randn((B, N, H * 32), device=dev, dtype=torch.bfloat16) for _ in range(4)]
qg, kg, vg = [t.clone().requires_grad_(True) for t in (q, k, v)]
sch = FlowMatchEulerDiscreteScheduler()
z = torch.randn((64, 1024, 768), device=dev, dtype=torch.bfloat16).requires_grad_(True)
y = torch.randn((64, 256, 768), device=dev, dtype=torch.bfloat16)
for _ in range(reps):
    flowmatch_noise(x0, sch, seed=1, offset=2, want_target=False)
    fused_denoise_loss(pred, x0, None, seed=1, offset=2)
    o = sana_linear_attention(qg, kg, vg, H)
    torch.autograd.grad(o, (qg, kg, vg), do)
    repa_cosine_loss(z, y, (32, 32))
torch.cuda.synchronize()
print("ok")
