#!/bin/bash
# linear-attention backward: tests (release build), timing, then the probe-build event trace
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_linattn_gpu.py tests/test_parity_k2_gpu.py -q -x 2>&1 | tail -2
YAT_PROBE_MASKS=7 python scripts/linattn_bwd_probe.py
YAT_NVCC_FLAGS="-DYAT_BWD_PROBE" python -m yat_b200.build --force > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
YAT_BWD_MASK=${YAT_BWD_MASK:-7} timeout 300 python scripts/linattn_bwd_trace.py 2>&1 | grep -A12 "epilogue\|issuer" | head -40
python -m yat_b200.build --force > gpurun_out/build.log 2>&1
