#!/bin/bash
# round 2, call e: baseline of the restored tree -- all -m gpu tests, smoke, bench, launch list
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi.txt
export YAT_REPORT_DIR=$PWD/gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --durations=10 > gpurun_out/t_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 30 gpurun_out/t_gpu.log | grep -v "^frame\|CUDAEvent"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -c 4000 gpurun_out/bench.log; grep -v "^frame\|CUDAEvent" gpurun_out/bench.err | tail -n 10
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-other-workloads --no-account --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "ncu rc=$?"
