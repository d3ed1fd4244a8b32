#!/bin/bash
# noise / loss iteration: parity tests, stand-alone kernel table, one full ncu capture of the two kernels
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || { cat gpurun_out/build.log; exit 1; }
timeout 900 python -m pytest tests/test_noise_gpu.py tests/test_loss_gpu.py tests/test_abi.py -q -m gpu 2>&1 | tail -12 | grep -v "^frame\|CUDAEvent"
timeout 300 python bench.py --kernels-only 2>&1 | tee gpurun_out/kernels_only.log | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(f\"{d['kernel']:22s} {d.get('us', 0):9.1f} us  {d.get('achieved', 0):7.0f} GB/s  frac {d.get('frac', 0):.3f}\" if 'us' in d else l[:200])
    else: print(l[:300], end='')
"
timeout 300 python scripts/prof_kernels.py 2 > gpurun_out/plain_prof.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"noise_bf16|loss_fused" -s 2 -c 2 \
    -o gpurun_out/prof_r01d -f python scripts/prof_kernels.py 2 > gpurun_out/ncu_prof.log 2>&1
echo "full capture rc=$?"; ls -la gpurun_out/*.ncu-rep | tail -3; tail -n 3 gpurun_out/ncu_prof.log
