"""GLUMBConv middle at the SANA-1.6B layer shape, forward + backward (for `ncu --set full`)."""
import sys
import torch
sys.path.insert(0, ".")
from yat_b200.glumb import glumb_mid
dev = torch.device("cuda")
B, H, W, Hd = 16, 32, 32, 5600
u = (torch.randn((B, H * W, 2 * Hd), device=dev) * 1.2).to(torch.bfloat16).requires_grad_(True)
w = (torch.randn((2 * Hd, 1, 3, 3), device=dev) * 0.3).to(torch.bfloat16).requires_grad_(True)
b = (torch.randn((2 * Hd,), device=dev) * 0.2).to(torch.bfloat16).requires_grad_(True)
dm = torch.randn((B, H * W, Hd), device=dev).to(torch.bfloat16)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    glumb_mid(u, w, b, H, W).backward(dm)
torch.cuda.synchronize()
print("ok")
