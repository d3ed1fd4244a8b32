"""dX GEMMs of the SANA block: torch's `dy @ W` (NN, what autograd runs) against `F.linear(dy, W^T stored row-major)`
(TN, the forward's kernel class), sustained (1 s of back-to-back launches per arm)."""
import json, sys, time
import torch

dev = torch.device("cuda:0")
M = 16384
shapes = {"qkv_dx": (6720, 2240), "to_out_dx": (2240, 2240), "conv_point_dx": (2240, 5600), "conv_inverted_dx": (11200, 2240)}


def sustained(fn, seconds=1.0):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n, t0 = 0, time.time()
    e0.record()
    while time.time() - t0 < seconds:
        for _ in range(50):
            fn()
        n += 50
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for name, (N, K) in shapes.items():          # forward: [M, K] x W[N, K]^T -> [M, N]; dX = dY[M, N] @ W[N, K] -> [M, K]
    dy = torch.randn((M, N), device=dev, dtype=torch.bfloat16)
    w = (torch.randn((N, K), device=dev) / K ** 0.5).to(torch.bfloat16)
    wt = w.t().contiguous()                  # [K, N] row-major
    fl = 2.0 * M * N * K
    r = {}
    for rep in range(2):
        us = sustained(lambda: dy @ w)
        r[f"nn_{rep}"] = (round(us, 1), round(fl / us / 1e6))
        us = sustained(lambda: torch.nn.functional.linear(dy, wt))
        r[f"tn_{rep}"] = (round(us, 1), round(fl / us / 1e6))
    print(name, json.dumps(r), flush=True)
