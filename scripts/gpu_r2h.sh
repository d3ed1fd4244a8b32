#!/bin/bash
# round 2: linear-attention backward (software-pipelined passes B / C): tests + stand-alone roofline lines
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 900 python -m pytest tests/test_linattn_gpu.py tests/test_parity_k2_gpu.py -q -x > gpurun_out/t_linattn.log 2>&1; echo "linattn pytest rc=$?"
tail -n 12 gpurun_out/t_linattn.log | grep -v "^frame\|CUDAEvent" | cut -c1-300
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
grep -o '"kernel": "\(glumb\|linattn\)[^}]*' gpurun_out/kernels_only.log | cut -c1-135
