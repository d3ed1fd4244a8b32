#!/bin/bash
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
timeout 300 python -m pytest tests/test_xattn_bwd_gpu.py tests/test_xattn_gpu.py -x -q 2>&1 | tail -15
timeout 300 python scripts/xattn_dev.py 2>&1 | tail -4
