"""GPU diagnostic: tcgen05 backward on one small case vs the CUDA-core restatement."""
import sys
import torch
sys.path.insert(0, ".")
from yat_b200 import _lib
dev = torch.device("cuda")
B, N, H = [int(a) for a in sys.argv[1:4]] if len(sys.argv) > 3 else (1, 256, 4)
g = torch.Generator(device=dev).manual_seed(1)
q, k, v, do = [torch.randn((B, N, H * 32), device=dev, generator=g).to(torch.bfloat16) for _ in range(4)]
outs = {}
for name in ("yat_linattn_bwd_simt", "yat_linattn_bwd"):
    dq, dk, dv = [torch.zeros_like(q) for _ in range(3)]
    rc = getattr(_lib.lib(), name)(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(do), B, N, H, 32,
                                   _lib.ptr(dq), _lib.ptr(dk), _lib.ptr(dv), _lib.current_stream())
    _lib.check(rc, name)
    torch.cuda.synchronize()
    outs[name] = (dq, dk, dv)
    print(name, "done", flush=True)
for a, b, n in zip(outs["yat_linattn_bwd"], outs["yat_linattn_bwd_simt"], ("dq", "dk", "dv")):
    print(n, "rel", ((a.float() - b.float()).norm() / b.float().norm()).item(), flush=True)
