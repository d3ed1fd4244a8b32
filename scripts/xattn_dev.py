"""Development harness of the cross-attention forward kernel (yat_xattn_fwd) at the SANA-1.6B layer size: timing with and without
a key-padding mask against the path the step runs (yat_head_pad + cuDNN SDPA with the bias folded into a head column), and -
in a YAT_BWD_PROBE build - the event trace of CTA 0 (role 0 = softmax thread 0, 1 = issuer)."""
import ctypes
import os
import sys
from collections import defaultdict

import torch

sys.path.insert(0, ".")
from yat_b200 import _lib  # noqa: E402
from yat_b200.attention import _sdpa_key_bias_folded  # noqa: E402

L = _lib.lib()
B, H, Nq, Lk, d = 16, 20, 1024, 320, 112
C = H * d
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
q = torch.randn((B, Nq, C), device=dev, generator=g).to(torch.bfloat16)
k, v = [torch.randn((B, Lk, C), device=dev, generator=g).to(torch.bfloat16) for _ in range(2)]
bias = torch.zeros((B, Lk), device=dev)
for b_, n_ in enumerate(torch.randint(20, 301, (B,), generator=torch.Generator().manual_seed(1)).tolist()):
    bias[b_, n_:] = -10000.0
o = torch.empty_like(q)
lse = torch.empty((B, H, Nq), device=dev)
st = torch.cuda.current_stream().cuda_stream


def run(masked):
    rc = L.yat_xattn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), bias.data_ptr() if masked else None, B, H, Nq, Lk, d, C, C,
                         d ** -0.5, o.data_ptr(), lse.data_ptr(), st)
    assert rc == 0, rc


def timeit(fn, iters=20, warm=3):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


bias4 = bias.view(B, 1, 1, Lk).to(torch.bfloat16)
with torch.no_grad():
    ref = _sdpa_key_bias_folded(q, k, v, bias4, H, d)
    run(True)
    rel = ((o.float() - ref.float()).norm() / ref.float().norm()).item()
    print(f"own masked {timeit(lambda: run(True)):.1f} us, own unmasked {timeit(lambda: run(False)):.1f} us, "
          f"library path {timeit(lambda: _sdpa_key_bias_folded(q, k, v, bias4, H, d)):.1f} us, rel diff vs library {rel:.2e}")

raw = ctypes.CDLL(os.path.join(os.path.dirname(_lib.__file__), "libyat_b200.so"))
if hasattr(raw, "yat_xattn_probe_read"):
    raw.yat_xattn_probe_read.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
    raw.yat_xattn_probe_reset()
    run(True)
    torch.cuda.synchronize()
    for role in range(4):
        buf = (ctypes.c_longlong * 4096)()
        n = raw.yat_xattn_probe_read(role, buf, 4096)
        ev = [(buf[i] >> 48, buf[i] & 0xffffffffffff) for i in range(n)]
        if not ev:
            continue
        print(f"== role {role}: {n} events, span {ev[-1][1] - ev[0][1]} cycles")
        acc = defaultdict(list)                     # codes >= 1000: per-item cycle accumulators (value / 16 in the low digits)
        for c0, _ in ev:
            if c0 >= 1000:
                acc[c0 // 1000].append((c0 % 1000) * 16)
        for kk in sorted(acc):
            print(f"   accumulator {kk}: n={len(acc[kk])} mean {sum(acc[kk]) / len(acc[kk]):8.0f} cycles per item")
        ev = [e for e in ev if e[0] < 1000]
        delta = defaultdict(list)
        for (c0, a), (c1, b) in zip(ev, ev[1:]):
            delta[(c0, c1)].append(b - a)
        for key in sorted(delta):
            dd = delta[key]
            print(f"   {key[0]:3d} -> {key[1]:3d}: n={len(dd):4d} mean {sum(dd) / len(dd):8.0f}  min {min(dd):7d}  max {max(dd):7d}")

# host cost of one C-ABI call (tensor-map encodes + launch), and the back-to-back device time bench.py reports
import time  # noqa: E402
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(200):
    run(True)
host_us = (time.perf_counter() - t0) / 200 * 1e6
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    run(True)
e1.record()
torch.cuda.synchronize()
print(f"host {host_us:.1f} us per call (queueing only); back to back {e0.elapsed_time(e1) * 20:.1f} us per call")

# backward: own (prep + dq + dkv) against the library path's backward (fwd + bwd of the folded SDPA minus its fwd)
do = torch.randn_like(q)
dq_, dk_, dv_ = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
ws = torch.empty(L.yat_xattn_bwd_workspace_bytes(B, H, Nq), dtype=torch.uint8, device=dev)
run(True)


def run_bwd(with_dkv=True):
    rc = L.yat_xattn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), bias.data_ptr(), o.data_ptr(), lse.data_ptr(), do.data_ptr(), B, H,
                         Nq, Lk, d, C, C, d ** -0.5, dq_.data_ptr(), dk_.data_ptr() if with_dkv else None,
                         dv_.data_ptr() if with_dkv else None, ws.data_ptr(), st)
    assert rc == 0, rc


qg, kg, vg = [t.clone().requires_grad_(True) for t in (q, k, v)]


def lib_fwd_bwd():
    y = _sdpa_key_bias_folded(qg, kg, vg, bias4, H, d)
    torch.autograd.grad(y, (qg, kg, vg), do)


with torch.no_grad():
    lib_f = timeit(lambda: _sdpa_key_bias_folded(q, k, v, bias4, H, d))
lib_fb = timeit(lib_fwd_bwd)
print(f"backward: own {timeit(run_bwd):.1f} us (prep + dq only {timeit(lambda: run_bwd(False)):.1f} us); "
      f"library fwd + bwd {lib_fb:.1f} us, fwd {lib_f:.1f} us -> bwd {lib_fb - lib_f:.1f} us")

# the same comparison at the power-capped clock the training step runs at: a GEMM burn before each arm, SM clock sampled by NVML
import pynvml  # noqa: E402
pynvml.nvmlInit()
hnd = pynvml.nvmlDeviceGetHandleByIndex(0)
ga = torch.randn((8192, 8192), device=dev).to(torch.bfloat16)
gb = torch.randn((8192, 8192), device=dev).to(torch.bfloat16)


def burn(seconds):
    t0 = time.time()
    while time.time() - t0 < seconds:
        for _ in range(20):
            ga @ gb
        torch.cuda.synchronize()


def capped(fn, n=150):
    burn(2.0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    mhz = pynvml.nvmlDeviceGetClockInfo(hnd, pynvml.NVML_CLOCK_SM)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3, mhz


with torch.no_grad():
    for rep in range(2):
        a_us, a_mhz = capped(lambda: run(True))
        b_us, b_mhz = capped(lambda: _sdpa_key_bias_folded(q, k, v, bias4, H, d))
        print(f"after a GEMM burn: own {a_us:.1f} us @ {a_mhz} MHz, library path {b_us:.1f} us @ {b_mhz} MHz")

# per-kernel durations of one backward call
from torch.profiler import ProfilerActivity, profile  # noqa: E402
run_bwd()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        run_bwd()
    torch.cuda.synchronize()
for ev in prof.key_averages():
    if ev.device_time_total > 0:
        print(f"   {ev.key[:70]:70s} {ev.device_time_total / ev.count:8.1f} us x{ev.count}")
