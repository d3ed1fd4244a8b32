#!/bin/bash
# glumb iteration: parity tests, stand-alone timing, then one ncu capture of the two kernels
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || { cat gpurun_out/build.log; exit 1; }
timeout 900 python -m pytest ${@:-tests/test_glumb_gpu.py} -q -m gpu -x 2>&1 | tail -8 | grep -v "^frame\|CUDAEvent"
timeout 300 python bench.py --kernels-only 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(f\"{d['kernel']:22s} {d.get('us', 0):9.1f} us  {d.get('achieved', 0):7.0f} GB/s  frac {d.get('frac', 0):.3f}\" if 'us' in d else l[:200])
    else: print(l[:300], end='')
"
timeout 300 python scripts/prof_glumb.py > gpurun_out/plain_glumb.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"glumb_fwd|glumb_bwd" -s 2 -c 2 \
    -o gpurun_out/prof_glumb -f python scripts/prof_glumb.py > gpurun_out/ncu_glumb.log 2>&1
echo "glumb capture rc=$?"; tail -n 2 gpurun_out/ncu_glumb.log
