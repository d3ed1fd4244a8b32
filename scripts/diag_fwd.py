"""GPU diagnostic: tcgen05 forward at growing item counts vs the CUDA-core restatement."""
import sys, time
import torch
sys.path.insert(0, ".")
from yat_b200.linear_attention import sana_linear_attention
what = sys.argv[1]
dev = torch.device("cuda")
if what == "noise":
    from yat_b200.noise import flowmatch_noise
    from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
    x0 = torch.randn((4096, 32, 32, 32), device=dev, dtype=torch.bfloat16)
    o = flowmatch_noise(x0, FlowMatchEulerDiscreteScheduler(), seed=1, offset=2, want_target=False)
    torch.cuda.synchronize(); print("noise ok", o.x_t.float().std().item())
elif what == "loss":
    from yat_b200.loss import fused_denoise_loss
    x0 = torch.randn((4096, 32, 32, 32), device=dev, dtype=torch.bfloat16)
    p = torch.randn_like(x0).requires_grad_(True)
    l = fused_denoise_loss(p, x0, None, seed=1, offset=2)
    torch.cuda.synchronize(); print("loss ok", l.item())
else:
    H, N = 70, 1024
    for B in [int(a) for a in sys.argv[2:]]:
        g = torch.Generator(device=dev).manual_seed(B)
        q, k, v = [torch.randn((B, N, H * 32), device=dev, generator=g).to(torch.bfloat16) for _ in range(3)]
        print("B", B, "items", B * 18, flush=True)
        ref = sana_linear_attention(q, k, v, H, impl="simt")
        torch.cuda.synchronize()
        t = time.time()
        o = sana_linear_attention(q, k, v, H, impl="sm100")
        torch.cuda.synchronize()
        rel = ((o.float() - ref.float()).norm() / ref.float().norm()).item()
        print("   rel", rel, "first-call ms", (time.time() - t) * 1e3, flush=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            o = sana_linear_attention(q, k, v, H, impl="sm100")
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("   us", ms * 1e3, "GB/s", 8 * B * N * H * 32 / ms / 1e6, flush=True)
