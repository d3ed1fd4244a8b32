#!/bin/bash
# cross-attention forward: tests + timing (release build), then the probe-build trace
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
timeout 300 python -m pytest tests/test_xattn_gpu.py -x -q 2>&1 | tail -8
timeout 200 python scripts/xattn_dev.py 2>&1 | tail -3
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
timeout 200 python scripts/xattn_dev.py 2>&1 | tail -40
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
