#!/bin/bash
# round 2, call f: GLUMB on the CTA-wide ring (tests + stand-alone roofline lines)
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python -m pytest tests/test_glumb_gpu.py -q -x > gpurun_out/t_glumb.log 2>&1; echo "glumb pytest rc=$?"
tail -n 8 gpurun_out/t_glumb.log | grep -v "^frame\|CUDAEvent"
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/kernels_only.log'):
    if l.startswith('{'):
        d = json.loads(l)
        for k, v in d.get('kernels', d).items():
            if isinstance(v, dict) and 'us' in v:
                print(f"{k:28s} {v['us']:9.1f} us  frac {v.get('frac', 0):.3f}")
PY
