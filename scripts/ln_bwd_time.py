"""Development timing: yat_ln_modulate_bwd(_res) alone at the K2 layer shape (CUDA events, back-to-back launches)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from yat_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
B, N, C = 16, 1024, 2240
x, dy, dres = [torch.randn((B, N, C), device=dev, dtype=torch.bfloat16) for _ in range(3)]
sc = torch.randn((B, C), device=dev, dtype=torch.bfloat16)
mean = x.float().mean(-1).reshape(-1).contiguous()
rstd = torch.rsqrt(x.float().var(-1, unbiased=False) + 1e-6).reshape(-1).contiguous()
dx = torch.empty_like(x)
dsc = torch.empty((B, C), device=dev, dtype=torch.bfloat16)
dsh = torch.empty_like(dsc)
ws = torch.empty(_lib.lib().yat_block_workspace_bytes(B, N, C), dtype=torch.uint8, device=dev)
L = _lib.lib()
st = _lib.current_stream()
for name, r in (("no_dres", None), ("dres", dres)):
    def fn():
        rc = L.yat_ln_modulate_bwd_res(_lib.ptr(x), _lib.ptr(dy), _lib.ptr(r), _lib.ptr(sc), C, _lib.ptr(mean), _lib.ptr(rstd),
                                       B, N, C, _lib.ptr(dx), _lib.ptr(dsc), _lib.ptr(dsh), _lib.ptr(ws), st)
        assert rc == 0
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 50 * 1e3
    nbytes = (8 if r is not None else 6) * B * N * C
    print(f"ln_modulate_bwd {name}: {us:.1f} us (incl. finish kernel)  {nbytes / us / 1e3:.0f} GB/s")
