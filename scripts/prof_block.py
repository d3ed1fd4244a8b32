"""Backward block ops + colsum at the SANA-1.6B layer shape (stand-alone CUDA-event times, or under ncu)."""
import sys
import torch
sys.path.insert(0, ".")
from yat_b200.block_ops import gate_residual, ln_modulate
from yat_b200.linear import colsum
dev = torch.device("cuda")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
B, N, C = 16, 1024, 2240
q, k, do = [torch.randn((B, N, C), device=dev, dtype=torch.bfloat16) for _ in range(3)]
mod = torch.randn((B, 6, C), device=dev, dtype=torch.bfloat16).requires_grad_(True)
sh, sc, gt = mod.chunk(6, dim=1)[:3]
big = torch.randn((B * N, 11200), device=dev, dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
xg = q.clone().requires_grad_(True)
yg = k.clone().requires_grad_(True)
y1 = ln_modulate(xg, sc, sh)
y2 = gate_residual(q, yg, gt)


def t(name, f, nbytes):
    f()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        f()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    us = sorted(ts)[len(ts) // 2]
    print(f"{name:20s} {us:8.1f} us  {nbytes / us / 1e3:7.0f} GB/s")


t("ln_modulate_bwd", lambda: torch.autograd.grad(y1, (xg, mod), do, retain_graph=True), 6 * B * N * C)
t("gate_residual_bwd", lambda: torch.autograd.grad(y2, (yg, mod), do, retain_graph=True), 6 * B * N * C)
t("colsum_2240", lambda: colsum(do.view(-1, C)), 2 * B * N * C)
t("colsum_11200", lambda: colsum(big), 2 * big.numel())
t("torch_sum_2240", lambda: do.view(-1, C).sum(0), 2 * B * N * C)
print("ok")
