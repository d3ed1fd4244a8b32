#!/bin/bash
# first GPU pass: per-file GPU tests (separate processes so one trapped kernel does not poison the rest)
mkdir -p gpurun_out
nvidia-smi > gpurun_out/smi.txt 2>&1
python -m yat_b200.build > gpurun_out/build.log 2>&1
for f in noise loss repa; do
  timeout 600 python -m pytest tests/test_${f}_gpu.py -q -m gpu > gpurun_out/test_${f}.log 2>&1
  echo "test_${f}: rc=$?" >> gpurun_out/summary.txt
done
timeout 600 python -m pytest tests/test_linattn_gpu.py -q -m gpu -k "simt" > gpurun_out/test_linattn_simt.log 2>&1
echo "test_linattn_simt: rc=$?" >> gpurun_out/summary.txt
timeout 300 python -m pytest tests/test_linattn_gpu.py -q -m gpu -k "forward_matches_oracle and sm100" > gpurun_out/test_linattn_sm100_fwd.log 2>&1
echo "test_linattn_sm100_fwd: rc=$?" >> gpurun_out/summary.txt
timeout 300 python -m pytest tests/test_step_gpu.py -q -m gpu -k "pad or forward" > gpurun_out/test_step_fwd.log 2>&1
echo "test_step_fwd: rc=$?" >> gpurun_out/summary.txt
timeout 900 python bench.py --steps 4 --warmup 3 --debug-linattn simt --no-cpu-baseline > gpurun_out/bench_debug_simt.log 2>&1
echo "bench_debug_simt: rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
tail -3 gpurun_out/test_*.log
tail -c 3000 gpurun_out/bench_debug_simt.log
