"""Development sweep: the own tcgen05 GEMM (tile width x ring depth) against the library call on the projection
shapes of one SANA-1.6B block at batch 16 (tokens 16384).  Writes gpurun_out/gemm_ab.json."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yat_b200.gemm import gemm_bf16_tn  # noqa: E402


def timeit(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def main():
    dev = torch.device("cuda:0")
    M = 16384
    shapes = {"qkv": (6720, 2240, False), "to_out": (2240, 2240, True), "conv_inverted": (11200, 2240, True),
              "conv_point": (2240, 5600, False), "caption_kv": (4480, 2240, True)}
    variants = [(224, 6), (256, 6), (224, 7), (256, 7), (224, 5), (256, 5), (224, 4)]
    res = {}
    for name, (N, K, has_bias) in shapes.items():
        m = M if name != "caption_kv" else 16 * 304
        a = torch.randn((m, K), device=dev, dtype=torch.bfloat16)
        w = (torch.randn((N, K), device=dev) / K ** 0.5).to(torch.bfloat16)
        b = torch.randn((N,), device=dev, dtype=torch.bfloat16) if has_bias else None
        out = torch.empty((m, N), device=dev, dtype=torch.bfloat16)
        fl = 2.0 * m * N * K
        r = {"shape": [m, N, K], "bias": has_bias}
        us = timeit(lambda: torch.nn.functional.linear(a, w, b))
        r["cublas"] = {"us": us, "tflops": fl / us / 1e6}
        for tn, st in variants:
            try:
                us = timeit(lambda: gemm_bf16_tn(a, w, b, tile_n=tn + 1000 * st, out=out))
                r[f"own_{tn}x{st}"] = {"us": us, "tflops": fl / us / 1e6, "vs_cublas": r["cublas"]["us"] / us}
            except Exception as e:  # noqa: BLE001
                r[f"own_{tn}x{st}"] = {"error": str(e)[:120]}
        res[name] = r
        print(name, json.dumps(r), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/gemm_ab.json", "w"), indent=1)


if __name__ == "__main__":
    main()
