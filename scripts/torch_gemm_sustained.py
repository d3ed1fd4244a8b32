"""Sustained (power-capped) time of torch's own GEMM call on the step's dominant shapes, for comparison with the per-algorithm
cuBLASLt sweep of scripts/lt_sweep.cu (same shapes, same operand distribution, 2 s pre-burn, then `seconds` per shape)."""
import sys
import time

import torch

seconds = float(sys.argv[1]) if len(sys.argv) > 1 else 0.8
dev = torch.device("cuda")
M, C = 16384, 2240
rnd = lambda *s: (torch.rand(s, device=dev) * 2 - 1).to(torch.bfloat16)  # noqa: E731
# (name, out_features): fwd y = x W^T; dgrad dx = dy W; wgrad dW = dy^T x
for name, cin, cout in (("qkv", C, 3 * C), ("conv_inverted", C, 5 * C), ("conv_point", 5 * C // 2, C), ("to_out", C, C)):
    x, w, dy = rnd(M, cin), rnd(cout, cin), rnd(M, cout)
    for kind, fn in (("fwd", lambda: torch.nn.functional.linear(x, w)), ("dgrad", lambda: dy @ w), ("wgrad", lambda: dy.t() @ x)):
        t0 = time.time()
        while time.time() - t0 < 2.0:
            for _ in range(50):
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
        us = e0.elapsed_time(e1) / n * 1e3
        print(f"torch {name} {kind}: {us:8.1f} us  {2.0 * M * cin * cout / us / 1e6:7.1f} TFLOP/s", flush=True)
