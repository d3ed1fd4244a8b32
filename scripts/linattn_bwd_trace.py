"""Event trace of CTA 0 of the linear-attention backward (YAT_BWD_PROBE build only): per role, the clock64 stamps of the
phase boundaries of pass B.  Prints the mean cycles between consecutive events of a tile, per role."""
import ctypes
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")
from yat_b200 import _lib  # noqa: E402

L = _lib.lib()
raw = ctypes.CDLL(os.path.join(os.path.dirname(_lib.__file__), "libyat_b200.so"))
raw.yat_bwd_probe_read.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
B, N, H, D = 16, 1024, 70, 32
dev = torch.device("cuda")
q, k, v, do = [torch.randn((B, N, H * D), device=dev, dtype=torch.bfloat16) for _ in range(4)]
dq, dk, dv = [torch.empty_like(q) for _ in range(3)]
st = torch.cuda.current_stream().cuda_stream


def run():
    rc = L.yat_linattn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), do.data_ptr(), B, N, H, D,
                           dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), st)
    assert rc == 0, rc


for _ in range(3):
    run()
torch.cuda.synchronize()
raw.yat_bwd_probe_reset()
run()
torch.cuda.synchronize()
names = {0: "epilogue", 1: "issuer", 2: "store", 3: "producer"}
for role in range(4):
    buf = (ctypes.c_longlong * 4096)()
    n = raw.yat_bwd_probe_read(role, buf, 4096)
    ev = [(buf[i] >> 48, buf[i] & 0xffffffffffff) for i in range(n)]
    if not ev:
        continue
    t0 = ev[0][1]
    print(f"== {names[role]}: {n} events, span {(ev[-1][1] - t0)} cycles")
    delta = defaultdict(list)
    for (c0, a), (c1, b) in zip(ev, ev[1:]):
        delta[(c0, c1)].append(b - a)
    for key in sorted(delta):
        d = delta[key]
        print(f"   {key[0]:3d} -> {key[1]:3d}: n={len(d):4d} mean {sum(d) / len(d):8.0f}  min {min(d):7d}  max {max(d):7d}")
