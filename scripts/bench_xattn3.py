"""Folded key-mask SDPA vs additive-mask SDPA with the real layouts (token-major projections), kernel breakdown."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from torch.profiler import ProfilerActivity, profile
from yat_b200.attention import _sdpa_key_bias_folded
dev = torch.device("cuda")
B, N, H, D, L = 16, 1024, 20, 112, 320
q0 = torch.randn(B, N, H * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
k0 = torch.randn(B, L, H * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
v0 = torch.randn(B, L, H * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
do = torch.randn(B, N, H * D, device=dev, dtype=torch.bfloat16)
m = (torch.rand(B, 1, 1, L, device=dev) < 0.5).to(torch.bfloat16)
bias = (1 - m) * -10000.0


def heads(t):
    return t.view(B, -1, H, D).transpose(1, 2)


def masked():
    o = F.scaled_dot_product_attention(heads(q0), heads(k0), heads(v0), attn_mask=bias)
    return o.transpose(1, 2).reshape(B, N, H * D)


def folded():
    return _sdpa_key_bias_folded(q0, k0, v0, bias, H, D)


for f, name in ((masked, "masked"), (folded, "folded")):
    for _ in range(3):
        f().backward(do)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    n = 20
    e[0].record()
    for _ in range(n):
        o = f()
    e[1].record()
    for _ in range(n):
        f().backward(do)
    e[2].record()
    torch.cuda.synchronize()
    print(f"{name}: fwd {e[0].elapsed_time(e[1]) / n * 1e3:.0f} us, fwd+bwd {e[1].elapsed_time(e[2]) / n * 1e3:.0f} us")
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        f().backward(do)
        torch.cuda.synchronize()
    for ev in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:12]:
        if ev.device_time_total > 0 and ev.device_type == torch.autograd.DeviceType.CUDA:
            print(f"    {ev.device_time_total:8.0f} us  x{ev.count:3d}  {ev.key[:110]}")
