#!/bin/bash
# round 2, full cycle: all -m gpu tests, smoke, bench, kernels-only, launch list, one full ncu capture of the own kernels
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi.txt
export YAT_REPORT_DIR=$PWD/gpurun_out
timeout 1500 python -m pytest tests -q -m gpu --durations=5 > gpurun_out/t_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 14 gpurun_out/t_gpu.log | grep -v "^frame\|CUDAEvent" | cut -c1-250
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log | cut -c1-200
timeout 900 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench.log; grep -v "^frame\|CUDAEvent" gpurun_out/bench.err | tail -n 5
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-other-workloads --no-account --no-cpu-baseline > gpurun_out/ncu.log 2>&1; echo "launch list rc=$?"
timeout 300 python scripts/prof_kernels.py 1 > gpurun_out/plain_prof.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"linattn_.*sm100|xattn_|qkv_linattn|gemm_bf16|noise_bf16|loss_fused|repa_|glumb_fwd|glumb_bwd|adamw|grad_sumsq|ln_modulate|gate_residual|column_pass|colsum_kernel|head_pad|head_unpad" -c 32 \
    -o gpurun_out/prof_r02g -f python scripts/prof_kernels.py 1 > gpurun_out/ncu_prof.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/*.ncu-rep | tail -2
