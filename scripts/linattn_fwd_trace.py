"""Event trace of CTA 0 of the linear-attention forward (YAT_BWD_PROBE build only): mean cycles between consecutive
phase-2 events per role (0 / 3 = epilogue groups 0 / 1, 1 = issuer, 2 = store)."""
import ctypes
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")
from yat_b200 import _lib  # noqa: E402

L = _lib.lib()
raw = ctypes.CDLL(os.path.join(os.path.dirname(_lib.__file__), "libyat_b200.so"))
raw.yat_fwd_probe_read.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
B, N, H, D = 16, 1024, 70, 32
dev = torch.device("cuda")
q, k, v = [torch.randn((B, N, H * D), device=dev, dtype=torch.bfloat16) for _ in range(3)]
o = torch.empty_like(q)
st = torch.cuda.current_stream().cuda_stream


def run():
    assert L.yat_linattn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), B, N, H, D, o.data_ptr(), st) == 0


for _ in range(3):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    run()
e1.record()
torch.cuda.synchronize()
print(f"forward: {e0.elapsed_time(e1) * 50:.1f} us back to back")
raw.yat_fwd_probe_reset()
run()
torch.cuda.synchronize()
for role in range(4):
    buf = (ctypes.c_longlong * 4096)()
    n = raw.yat_fwd_probe_read(role, buf, 4096)
    ev = [(buf[i] >> 48, buf[i] & 0xffffffffffff) for i in range(n)]
    if not ev:
        continue
    print(f"== role {role}: {n} events, span {ev[-1][1] - ev[0][1]} cycles")
    delta = defaultdict(list)
    for (c0, a), (c1, b) in zip(ev, ev[1:]):
        delta[(c0, c1)].append(b - a)
    for key in sorted(delta):
        d = delta[key]
        print(f"   {key[0]:3d} -> {key[1]:3d}: n={len(d):4d} mean {sum(d) / len(d):8.0f}  min {min(d):7d}  max {max(d):7d}")
