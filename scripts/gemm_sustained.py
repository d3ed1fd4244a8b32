"""Sustained (power-capped) throughput of the own tcgen05 GEMM vs the library call: ~1.5 s of back-to-back launches
per arm, alternating, with the SM clock and board power sampled through NVML while each arm runs.  The training step
runs its GEMMs in exactly this regime (sw_power_cap for the whole step), unlike the 30-launch burst of gemm_ab.py."""
import json
import os
import sys
import threading
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yat_b200.gemm import gemm_bf16_tn  # noqa: E402

import pynvml  # noqa: E402

pynvml.nvmlInit()
H = pynvml.nvmlDeviceGetHandleByIndex(0)


class Sampler:
    def __enter__(self):
        self.stop, self.mhz, self.w = False, [], []
        self.t = threading.Thread(target=self.run)
        self.t.start()
        return self

    def run(self):
        while not self.stop:
            self.mhz.append(pynvml.nvmlDeviceGetClockInfo(H, pynvml.NVML_CLOCK_SM))
            self.w.append(pynvml.nvmlDeviceGetPowerUsage(H) / 1000.0)
            time.sleep(0.02)

    def __exit__(self, *a):
        self.stop = True
        self.t.join()


def sustained(fn, seconds=1.5):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    with Sampler() as s:
        t0 = time.time()
        e0.record()
        while time.time() - t0 < seconds:
            for _ in range(50):
                fn()
            n += 50
            torch.cuda.synchronize()
        e1.record()
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    k = len(s.mhz) // 3
    return us, sorted(s.mhz[k:])[len(s.mhz[k:]) // 2], sorted(s.w[k:])[len(s.w[k:]) // 2]


def main():
    dev = torch.device("cuda:0")
    M = 16384
    shapes = {"qkv": (6720, 2240), "conv_inverted": (11200, 2240), "conv_point": (2240, 5600), "to_out": (2240, 2240)}
    res = {}
    for name, (N, K) in shapes.items():
        a = torch.randn((M, K), device=dev, dtype=torch.bfloat16)
        w = (torch.randn((N, K), device=dev) / K ** 0.5).to(torch.bfloat16)
        out = torch.empty((M, N), device=dev, dtype=torch.bfloat16)
        fl = 2.0 * M * N * K
        r = {}
        for rep in range(2):
            for arm, fn in (("cublas", lambda: torch.nn.functional.linear(a, w)), ("own", lambda: gemm_bf16_tn(a, w, out=out))):
                us, mhz, watts = sustained(fn)
                r[f"{arm}_{rep}"] = {"us": round(us, 1), "tflops": round(fl / us / 1e6), "sm_mhz": mhz, "watts": round(watts),
                                     "gflop_per_joule": round(fl / us / 1e3 / watts)}
        res[name] = r
        print(name, json.dumps(r), flush=True)
    json.dump(res, open("gpurun_out/gemm_sustained.json", "w"), indent=1)


if __name__ == "__main__":
    main()
