#!/bin/bash
# round 2, call a: all -m gpu tests (reports into gpurun_out/), smoke, the bench with accounting + other workloads
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi.txt
export YAT_REPORT_DIR=$PWD/gpurun_out
timeout 1700 python -m pytest tests -q -m gpu --durations=15 > gpurun_out/t_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 40 gpurun_out/t_gpu.log | grep -v "^frame\|CUDAEvent"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 1200 python bench.py --steps 8 --warmup 3 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/bench.log; grep -v "^frame\|CUDAEvent" gpurun_out/bench.err | tail -n 15
