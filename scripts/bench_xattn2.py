"""cuDNN SDPA variants for the SANA cross-attention shape: additive mask vs no mask, d = 112 vs 128."""
import torch
import torch.nn.functional as F
dev = torch.device("cuda")
B, N, H, L = 16, 1024, 20, 320


def run(D, masked, name, Lk=L):
    q = torch.randn(B, H, N, D, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(B, H, Lk, D, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(B, H, Lk, D, device=dev, dtype=torch.bfloat16, requires_grad=True)
    do = torch.randn(B, H, N, D, device=dev, dtype=torch.bfloat16)
    bias = None
    if masked:
        m = (torch.rand(B, 1, 1, Lk, device=dev) < 0.5).to(torch.bfloat16)
        bias = (1 - m) * -10000.0
    f = lambda: F.scaled_dot_product_attention(q, k, v, attn_mask=bias)
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


run(112, True, "d112 additive mask")
run(112, False, "d112 no mask")
run(128, False, "d128 no mask")
run(128, True, "d128 additive mask")
run(128, False, "d128 no mask L=192", 192)
run(112, False, "d112 no mask L=192", 192)
