#!/bin/bash
# experiment: sustained throughput of every cuBLASLt heuristic candidate on the step's dominant GEMM shapes (torch's bundled
# library), next to torch's own call on the same shapes
mkdir -p gpurun_out
nvcc -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o /tmp/lt_sweep scripts/lt_sweep.cu -lcublasLt || exit 1
export LD_LIBRARY_PATH=/opt/prime-rl/.venv/lib/python3.12/site-packages/nvidia/cublas/lib:$LD_LIBRARY_PATH
timeout 900 /tmp/lt_sweep ${1:-0.8} > gpurun_out/lt_sweep.txt 2>&1; echo "rc=$?"
timeout 600 python scripts/torch_gemm_sustained.py ${1:-0.8} > gpurun_out/torch_gemm_sustained.txt 2>&1; echo "rc=$?"
python - <<'PY'
import re
best = {}
cur = None
for line in open("gpurun_out/lt_sweep.txt"):
    m = re.match(r"== (\S+ \S+) ", line)
    if m:
        cur = m.group(1); best[cur] = []
        continue
    m = re.search(r"#(\d+)\s+id=(\d+)\s+tile=(\d+).*?waves=([\d.]+)\s+([\d.]+) us", line)
    if m and cur:
        best[cur].append((float(m.group(5)), int(m.group(1)), int(m.group(3))))
torch_us = {}
for line in open("gpurun_out/torch_gemm_sustained.txt"):
    m = re.match(r"torch (\S+ \S+):\s+([\d.]+) us", line)
    if m:
        torch_us[m.group(1)] = float(m.group(2))
for k, v in best.items():
    if not v: continue
    b = min(v); first = [x for x in v if x[1] == 0][0]
    print(f"{k:22s} torch {torch_us.get(k, 0):7.1f}  lt#0 {first[0]:7.1f} (tile {first[2]})  best {b[0]:7.1f} (#{b[1]} tile {b[2]})  torch/best {torch_us.get(k, 0) / b[0]:.3f}")
PY
