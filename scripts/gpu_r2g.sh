#!/bin/bash
# round 2, call g: GLUMB tests, timings, then one full ncu capture of the two kernels
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_glumb_gpu.py -q > gpurun_out/t_glumb.log 2>&1; echo "glumb pytest rc=$?"
tail -n 4 gpurun_out/t_glumb.log | grep -v "^frame\|CUDAEvent"
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
grep -o '"kernel": "glumb[^}]*' gpurun_out/kernels_only.log | cut -c1-260
timeout 300 python scripts/prof_glumb.py > gpurun_out/plain_prof_glumb.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"glumb_fwd|glumb_bwd" -c 2 \
    -o gpurun_out/prof_glumb_r02 -f python scripts/prof_glumb.py > gpurun_out/ncu_prof.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/*.ncu-rep | tail -3
