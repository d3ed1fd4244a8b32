#!/bin/bash
# experiment: event trace of the linear-attention backward (probe build, NOT the product library)
mkdir -p gpurun_out
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
YAT_BWD_MASK=${YAT_BWD_MASK:-7} timeout 300 python scripts/linattn_bwd_trace.py 2>&1 | tail -60
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
