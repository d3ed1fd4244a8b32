#!/bin/bash
# launch list of one bench step + full captures of the hot-path kernels (each after a plain run exited 0)
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_step_gpu.py -q -m gpu > gpurun_out/t_step.log 2>&1; echo "step rc=$?"; tail -n 3 gpurun_out/t_step.log
python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"; wc -l gpurun_out/launches.csv
python scripts/prof_kernels.py 2 > gpurun_out/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"linattn_.*sm100|noise_bf16|loss_fused" -s 4 -c 4 \
    -o gpurun_out/prof_r01 -f python scripts/prof_kernels.py 2 > gpurun_out/ncu_prof.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/*.ncu-rep; tail -n 5 gpurun_out/ncu_prof.log
