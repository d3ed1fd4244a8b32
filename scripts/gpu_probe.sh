#!/bin/bash
# experiment: per-pass timing of the linear-attention backward (probe build, NOT the product library)
mkdir -p gpurun_out
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1
YAT_PROBE_MASKS=${YAT_PROBE_MASKS:-7,1,2,4} timeout 600 python scripts/linattn_bwd_probe.py 2>&1 | tail -8
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
