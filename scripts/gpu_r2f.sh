#!/bin/bash
# round 2: GLUMB tests + stand-alone roofline lines
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_glumb_gpu.py -q > gpurun_out/t_glumb.log 2>&1; echo "glumb pytest rc=$?"
tail -n 8 gpurun_out/t_glumb.log | grep -v "^frame\|CUDAEvent" | cut -c1-300
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
grep -o '"kernel": "glumb[^}]*' gpurun_out/kernels_only.log | cut -c1-130
