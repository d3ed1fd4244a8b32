"""One launch of every hot-path kernel at its roofline shape (for `ncu --set full`)."""
import sys
import torch
sys.path.insert(0, ".")
from yat_b200.block_ops import gate_residual, ln_modulate
from yat_b200.attention import _HeadPad, _HeadUnpad
from yat_b200.glumb import glumb_mid
from yat_b200.linear import colsum
from yat_b200.linear_attention import sana_linear_attention, sana_linear_attention_proj
from yat_b200.gemm import gemm_bf16_tn
from yat_b200.loss import fused_denoise_loss
from yat_b200.noise import flowmatch_noise
from yat_b200.optim import FusedClipAdamW
from yat_b200.repa import repa_cosine_loss
from yat_b200.schedulers import FlowMatchEulerDiscreteScheduler
dev = torch.device("cuda")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
x0 = torch.randn((4096, 32, 32, 32), device=dev, dtype=torch.bfloat16)
pred = torch.randn_like(x0).requires_grad_(True)
B, N, H = 16, 1024, 70
C = H * 32
q, k, v, do = [torch.randn((B, N, C), device=dev, dtype=torch.bfloat16) for _ in range(4)]
qg, kg, vg = [t.clone().requires_grad_(True) for t in (q, k, v)]
sch = FlowMatchEulerDiscreteScheduler()
z = torch.randn((64, 1024, 768), device=dev, dtype=torch.bfloat16).requires_grad_(True)
y = torch.randn((64, 256, 768), device=dev, dtype=torch.bfloat16)
mod = torch.randn((B, 6, C), device=dev, dtype=torch.bfloat16).requires_grad_(True)
sh, sc, gt = mod.chunk(6, dim=1)[:3]
Hd = 5600
u = (torch.randn((B, N, 2 * Hd), device=dev) * 1.2).to(torch.bfloat16).requires_grad_(True)
wd = (torch.randn((2 * Hd, 1, 3, 3), device=dev) * 0.3).to(torch.bfloat16).requires_grad_(True)
bd = (torch.randn((2 * Hd,), device=dev) * 0.2).to(torch.bfloat16).requires_grad_(True)
dm = torch.randn((B, N, Hd), device=dev).to(torch.bfloat16)
params = [torch.nn.Parameter(torch.randn(2240 * 2240, device=dev, dtype=torch.bfloat16)) for _ in range(64)]   # 321 M
opt = FusedClipAdamW(params, lr=1e-5, weight_decay=0.0)
for p in params:
    p.grad = torch.randn_like(p) * 1e-3
wq = (torch.randn((3 * C, C), device=dev) / C ** 0.5).to(torch.bfloat16)
xk, xv = [torch.randn((B, 320, C), device=dev).to(torch.bfloat16) for _ in range(2)]     # cross-attention: 20 heads x 112, 320 keys
xo, xlse = torch.empty_like(q), torch.empty((B, 20, N), device=dev)
xdq, xdk, xdv = torch.empty_like(q), torch.empty_like(xk), torch.empty_like(xv)
from yat_b200 import _lib as _L  # noqa: E402
for _ in range(reps):
    _L.lib().yat_xattn_fwd(q.data_ptr(), xk.data_ptr(), xv.data_ptr(), None, B, 20, N, 320, 112, C, C, 112 ** -0.5, xo.data_ptr(),
                           xlse.data_ptr(), torch.cuda.current_stream().cuda_stream)
    _xws = torch.empty(_L.lib().yat_xattn_bwd_workspace_bytes(B, 20, N), dtype=torch.uint8, device=dev)
    _L.lib().yat_xattn_bwd(q.data_ptr(), xk.data_ptr(), xv.data_ptr(), None, xo.data_ptr(), xlse.data_ptr(), do.data_ptr(), B, 20, N, 320, 112,
                           C, C, 112 ** -0.5, xdq.data_ptr(), xdk.data_ptr(), xdv.data_ptr(), _xws.data_ptr(), torch.cuda.current_stream().cuda_stream)
    with torch.no_grad():
        sana_linear_attention_proj(q, wq, H)          # fused projection + phase 1, then the state-mode attention forward
        gemm_bf16_tn(q.view(B * N, C), wq)            # the plain own GEMM of the same projection
    flowmatch_noise(x0, sch, seed=1, offset=2, want_target=False)
    fused_denoise_loss(pred, x0, None, seed=1, offset=2)
    o = sana_linear_attention(qg, kg, vg, H)
    torch.autograd.grad(o, (qg, kg, vg), do)
    repa_cosine_loss(z, y, (32, 32))
    xg = q.clone().requires_grad_(True)
    torch.autograd.grad(ln_modulate(xg, sc, sh), (xg, mod), do)
    yg = k.clone().requires_grad_(True)
    torch.autograd.grad(gate_residual(q, yg, gt), (yg, mod), do)
    glumb_mid(u, wd, bd, 32, 32).backward(dm)
    colsum(do.view(-1, C))
    _HeadUnpad.apply(_HeadPad.apply(q, 20, 112, 128, 1.0, None), 112)
    opt.step(max_norm=1.0)
torch.cuda.synchronize()
print("ok")
