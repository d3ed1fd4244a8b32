"""cuDNN SDPA with d_qk = 128 (mask column folded in) and d_v = 112: does the library take different q/k and v head dims?"""
import torch
import torch.nn.functional as F
dev = torch.device("cuda")
B, N, H, L = 16, 1024, 20, 320


def run(dqk, dv, name):
    q = torch.randn(B, N, H, dqk, device=dev, dtype=torch.bfloat16, requires_grad=True)
    k = torch.randn(B, L, H, dqk, device=dev, dtype=torch.bfloat16, requires_grad=True)
    v = torch.randn(B, L, H, dv, device=dev, dtype=torch.bfloat16, requires_grad=True)
    do = torch.randn(B, N, H, dv, device=dev, dtype=torch.bfloat16)
    f = lambda: F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), scale=112 ** -0.5)
    o = f()
    print(name, "out", tuple(o.shape), "strides", o.stride())
    for _ in range(3):
        f().backward(do.transpose(1, 2))
    torch.cuda.synchronize()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    n = 20
    e[0].record()
    for _ in range(n):
        f()
    e[1].record()
    for _ in range(n):
        f().backward(do.transpose(1, 2))
    e[2].record()
    torch.cuda.synchronize()
    print(f"{name}: fwd {e[0].elapsed_time(e[1]) / n * 1e3:.0f} us, fwd+bwd {e[1].elapsed_time(e[2]) / n * 1e3:.0f} us")


run(128, 128, "qk128 v128")
run(128, 112, "qk128 v112")
