#!/bin/bash
# One GPU call: all -m gpu tests, smoke, the bench, then (only after each plain run exited 0)
# the ncu launch list of one bench step and one full capture of the hot-path kernels.
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi.txt
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/t_gpu.log 2>&1; echo "pytest rc=$?"
tail -n 15 gpurun_out/t_gpu.log | grep -v "^frame\|CUDAEvent"
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 8 --warmup 3 > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -c 5000 gpurun_out/bench.log; grep -v "^frame\|CUDAEvent" gpurun_out/bench.err | tail -n 15
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
timeout 300 python scripts/prof_kernels.py 1 > gpurun_out/plain_prof.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"linattn_.*sm100|noise_bf16|loss_fused|repa_|glumb_fwd|glumb_bwd|adamw|grad_sumsq|ln_modulate|gate_residual|column_pass|colsum_kernel|head_pad|head_unpad" -c 19 \
    -o gpurun_out/prof_r01h -f python scripts/prof_kernels.py 1 > gpurun_out/ncu_prof.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/*.ncu-rep; tail -n 5 gpurun_out/ncu_prof.log
