#!/bin/bash
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 300 python -m pytest tests/test_block_ops_gpu.py -q -m gpu -x 2>&1 | tail -5
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
