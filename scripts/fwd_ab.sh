python -m yat_b200.build --force > /dev/null
python - <<'PY'
import torch, sys
sys.path.insert(0, '.')
from yat_b200 import _lib
L = _lib.lib()
B, N, H, D = 16, 1024, 70, 32
dev = torch.device('cuda')
q, k, v = [torch.randn((B, N, H * D), device=dev, dtype=torch.bfloat16) for _ in range(3)]
o = torch.empty_like(q)
st = torch.cuda.current_stream().cuda_stream
def run(): assert L.yat_linattn_fwd(q.data_ptr(), k.data_ptr(), v.data_ptr(), B, N, H, D, o.data_ptr(), st) == 0
for _ in range(5): run()
torch.cuda.synchronize()
res = []
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): run()
    e1.record(); torch.cuda.synchronize()
    res.append(e0.elapsed_time(e1) * 50)
print("fwd us:", [round(r, 1) for r in res])
PY
