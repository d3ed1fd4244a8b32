#!/bin/bash
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || { cat gpurun_out/build.log; exit 1; }
timeout 600 python -m pytest tests/test_block_ops_gpu.py -q -m gpu -x 2>&1 | tail -5 | grep -v "^frame\|CUDAEvent"
timeout 300 python scripts/prof_block.py 1 > gpurun_out/plain_block.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"ln_modulate_bwd_fused|colsum_kernel|column_pass" -c 6 \
    -o gpurun_out/prof_block -f python scripts/prof_block.py 1 > gpurun_out/ncu_block.log 2>&1
echo "capture rc=$?"; tail -n 2 gpurun_out/ncu_block.log
