#!/bin/bash
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 300 python -m pytest tests/test_linattn_gpu.py -q -m gpu -x -k "sm100 and (backward_matches or golden)" > gpurun_out/t_bwd_small.log 2>&1; echo "bwd_small rc=$?"
tail -n 25 gpurun_out/t_bwd_small.log | grep -v "^frame\|CUDAEvent"
timeout 600 python -m pytest tests/test_linattn_gpu.py -q -m gpu > gpurun_out/t_linattn.log 2>&1; echo "linattn rc=$?"
tail -n 30 gpurun_out/t_linattn.log | grep -v "^frame\|CUDAEvent"
timeout 600 python -m pytest tests/test_step_gpu.py -q -m gpu > gpurun_out/t_step.log 2>&1; echo "step rc=$?"
tail -n 30 gpurun_out/t_step.log | grep -v "^frame\|CUDAEvent"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -c 6000 gpurun_out/bench.log; grep -v "^frame\|CUDAEvent" gpurun_out/bench.err | tail -n 15
