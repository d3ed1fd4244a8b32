#!/bin/bash
mkdir -p gpurun_out
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
timeout 200 python scripts/xattn_bwd_trace.py 2>&1 | tail -30
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
