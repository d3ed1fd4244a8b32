"""Timing experiment: linear-attention backward with subsets of its three passes (YAT_BWD_PROBE build only; the
results of a partial run are wrong by construction).  Prints us per launch for each pass mask."""
import os, subprocess, sys
sys.path.insert(0, ".")
if len(sys.argv) > 1:
    import torch
    from yat_b200 import _lib
    L = _lib.lib()
    B, N, H, D = 16, 1024, 70, 32
    dev = torch.device("cuda")
    q, k, v, do = [torch.randn((B, N, H * D), device=dev, dtype=torch.bfloat16) for _ in range(4)]
    dq, dk, dv = [torch.empty_like(q) for _ in range(3)]
    st = torch.cuda.current_stream().cuda_stream
    def run():
        rc = L.yat_linattn_bwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), do.data_ptr(), B, N, H, D,
                               dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), st)
        assert rc == 0, rc
    for _ in range(5):
        run()
    big = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for _ in range(20):
        big.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); run(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        run()
    e1.record(); torch.cuda.synchronize()
    print(f"mask {os.environ.get('YAT_BWD_MASK', '7')}: median {ts[len(ts) // 2]:.1f} us  best {ts[0]:.1f} us  "
          f"(L2 flushed between launches);  back to back {e0.elapsed_time(e1) * 1e3 / 20:.1f} us")
else:
    for m in [int(x) for x in os.environ.get('YAT_PROBE_MASKS', '7,1,2,4,3,6').split(',')]:
        env = dict(os.environ, YAT_BWD_MASK=str(m))
        subprocess.run([sys.executable, __file__, "run"], env=env, check=False)
