"""Cross-attention of a SANA-1.6B block: torch SDPA (cuDNN, padded + mask) vs flash_attn varlen (library) fwd+bwd."""
import sys
import torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from flash_attn import flash_attn_varlen_func
dev = torch.device("cuda")
B, N, H, D = 16, 1024, 20, 112
g = torch.Generator().manual_seed(0)
lens = [int(torch.randint(20, 301, (1,), generator=g)) for _ in range(B)]
Lmax = (max(lens) + 63) // 64 * 64
print("lens", lens, "Lmax", Lmax, "sum", sum(lens))
q = torch.randn(B, N, H * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
k = torch.randn(B, Lmax, H * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
v = torch.randn(B, Lmax, H * D, device=dev, dtype=torch.bfloat16, requires_grad=True)
mask = torch.zeros(B, Lmax, device=dev)
for i, l in enumerate(lens):
    mask[i, :l] = 1
bias = ((1 - mask) * -10000.0).to(torch.bfloat16).view(B, 1, 1, Lmax)
do = torch.randn(B, N, H * D, device=dev, dtype=torch.bfloat16)


def sdpa():
    o = F.scaled_dot_product_attention(q.view(B, N, H, D).transpose(1, 2), k.view(B, Lmax, H, D).transpose(1, 2),
                                       v.view(B, Lmax, H, D).transpose(1, 2), attn_mask=bias)
    return o.transpose(1, 2).reshape(B, N, H * D)


idx = torch.cat([torch.arange(l) + i * Lmax for i, l in enumerate(lens)]).to(dev)
cu_k = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=dev)
cu_q = torch.arange(0, (B + 1) * N, N, dtype=torch.int32, device=dev)


def fa():
    kp = k.view(B * Lmax, H, D).index_select(0, idx)
    vp = v.view(B * Lmax, H, D).index_select(0, idx)
    o = flash_attn_varlen_func(q.view(B * N, H, D), kp, vp, cu_q, cu_k, N, max(lens))
    return o.view(B, N, H * D)


def timeit(f, name):
    for _ in range(3):
        o = f()
        o.backward(do)
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    n = 20
    e[0].record()
    for _ in range(n):
        o = f()
    e[1].record()
    for _ in range(n):
        o = f()
        o.backward(do)
    e[2].record()
    torch.cuda.synchronize()
    fw = e[0].elapsed_time(e[1]) / n * 1e3
    fb = e[1].elapsed_time(e[2]) / n * 1e3
    print(f"{name}: fwd {fw:.0f} us, fwd+bwd {fb:.0f} us")
    return o


o1 = timeit(sdpa, "sdpa")
g1 = [t.grad.clone() for t in (q, k, v)]
for t in (q, k, v):
    t.grad = None
o2 = timeit(fa, "flash_varlen")
o1 = sdpa(); o1.backward(do); g1 = [t.grad.clone() for t in (q, k, v)]
for t in (q, k, v):
    t.grad = None
o2 = fa(); o2.backward(do); g2 = [t.grad.clone() for t in (q, k, v)]
rel = lambda a, b: ((a.float() - b.float()).norm() / b.float().norm()).item()
m3 = mask.view(B, Lmax, 1)
print("out rel", rel(o2, o1), "dq", rel(g2[0], g1[0]), "dk", rel(g2[1] * m3, g1[1] * m3), "dv", rel(g2[2] * m3, g1[2] * m3))
