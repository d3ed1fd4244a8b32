#!/bin/bash
# linear-attention forward (16 warps, Ks on the tensor core): tests (release build), timing, then the probe-build trace
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_linattn_gpu.py tests/test_qkv_linattn_gpu.py tests/test_parity_k2_gpu.py -q -x 2>&1 | tail -3
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
timeout 300 python scripts/linattn_fwd_trace.py 2>&1 | head -60
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
