#!/bin/bash
# round 2: linear-attention backward tests + timing, then the per-pass probe build
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 900 python -m pytest tests/test_linattn_gpu.py tests/test_parity_k2_gpu.py -q -x > gpurun_out/t_linattn.log 2>&1; echo "linattn pytest rc=$?"
tail -n 6 gpurun_out/t_linattn.log | grep -v "^frame\|CUDAEvent" | cut -c1-300
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1
YAT_PROBE_MASKS=${YAT_PROBE_MASKS:-7,1,2,4} timeout 600 python scripts/linattn_bwd_probe.py 2>&1 | tail -8
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
