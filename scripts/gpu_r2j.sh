#!/bin/bash
# round 2: fused projection + linear-attention phase 1 (yat_qkv_linattn_fwd / yat_linattn_fwd_state)
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 300 python -m pytest tests/test_qkv_linattn_gpu.py tests/test_linattn_gpu.py -q -x > gpurun_out/t_qkv.log 2>&1; echo "qkv pytest rc=$?"
grep -v "^frame\|CUDAEvent" gpurun_out/t_qkv.log | tail -n 30 | cut -c1-400
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
grep -o '"kernel": "linattn[^}]*' gpurun_out/kernels_only.log | cut -c1-140
