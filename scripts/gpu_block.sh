#!/bin/bash
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || { cat gpurun_out/build.log; exit 1; }
timeout 600 python -m pytest tests/test_block_ops_gpu.py -q -m gpu -x 2>&1 | tail -5 | grep -v "^frame\|CUDAEvent"
timeout 300 python scripts/prof_block.py 1 > gpurun_out/plain_block.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"ln_modulate_bwd" -c 4 python scripts/prof_block.py 1 2>&1 | grep -E "ln_modulate_bwd|duration|inst_executed|issue_active|warps_active"
YAT_LN_BWD_PAIRS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"ln_modulate_bwd" -c 2 python scripts/prof_block.py 1 2>&1 | grep -E "ln_modulate_bwd|duration"
