"""Event trace of CTA 0 of the cross-attention backward kernels (YAT_BWD_PROBE build only)."""
import ctypes
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")
from yat_b200 import _lib  # noqa: E402

L = _lib.lib()
raw = ctypes.CDLL(os.path.join(os.path.dirname(_lib.__file__), "libyat_b200.so"))
raw.yat_xbwd_probe_read.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
B, H, Nq, Lk, d = 16, 20, 1024, 320, 112
C = H * d
dev = torch.device("cuda")
q, do = [torch.randn((B, Nq, C), device=dev).to(torch.bfloat16) for _ in range(2)]
k, v = [torch.randn((B, Lk, C), device=dev).to(torch.bfloat16) for _ in range(2)]
o, dq = torch.empty_like(q), torch.empty_like(q)
dk, dv = torch.empty_like(k), torch.empty_like(v)
lse = torch.empty((B, H, Nq), device=dev)
ws = torch.empty(L.yat_xattn_bwd_workspace_bytes(B, H, Nq), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
assert L.yat_xattn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), None, B, H, Nq, Lk, d, C, C, d ** -0.5, o.data_ptr(), lse.data_ptr(), st) == 0


def run():
    assert L.yat_xattn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), None, o.data_ptr(), lse.data_ptr(), do.data_ptr(), B, H, Nq, Lk, d,
                           C, C, d ** -0.5, dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), ws.data_ptr(), st) == 0


for _ in range(3):
    run()
torch.cuda.synchronize()
raw.yat_xbwd_probe_reset()
run()
torch.cuda.synchronize()
for role in range(4):
    buf = (ctypes.c_longlong * 4096)()
    n = raw.yat_xbwd_probe_read(role, buf, 4096)
    ev = [(buf[i] >> 48, buf[i] & 0xffffffffffff) for i in range(n)]
    if not ev:
        continue
    print(f"== role {role}: {n} events, span {ev[-1][1] - ev[0][1]} cycles")
    acc = defaultdict(list)
    for c0, _ in ev:
        if c0 >= 1000:
            acc[c0 // 1000].append((c0 % 1000) * 16)
    for kk in sorted(acc):
        print(f"   accumulator {kk}: n={len(acc[kk])} mean {sum(acc[kk]) / len(acc[kk]):8.0f} cycles")
    ev = [e for e in ev if e[0] < 1000]
    delta = defaultdict(list)
    for (c0, a), (c1, b) in zip(ev, ev[1:]):
        delta[(c0, c1)].append(b - a)
    for key in sorted(delta):
        dd = delta[key]
        print(f"   {key[0]:3d} -> {key[1]:3d}: n={len(dd):4d} mean {sum(dd) / len(dd):8.0f}  min {min(dd):7d}  max {max(dd):7d}")
