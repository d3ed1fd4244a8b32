#!/bin/bash
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || tail -20 gpurun_out/build.log
timeout 600 python -m pytest tests/test_xattn_gpu.py tests/test_step_gpu.py tests/test_hooks_gpu.py tests/test_block_ops_gpu.py -x -q 2>&1 | tail -8
timeout 200 python scripts/xattn_dev.py 2>&1 | tail -2
