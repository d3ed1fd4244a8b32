#!/bin/bash
# round 2, call d: LN backward on the shared-memory row ring
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_block_ops_gpu.py -q -x > gpurun_out/t_block.log 2>&1; echo "block pytest rc=$?"
tail -n 8 gpurun_out/t_block.log | grep -v "^frame\|CUDAEvent"
timeout 300 python scripts/ln_bwd_time.py 2>&1 | tail -4
