#!/bin/bash
# round 2: kernels-only lines (fused projection + attention A/B among them)
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/kernels_only.log'):
    if l.startswith('{'):
        d = json.loads(l)
        print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in d.items() if k in ('kernel', 'us', 'frac', 'achieved', 'unit', 'vs_cublas_time', 'fused_us', 'unfused_us', 'fused_over_unfused_time', 'error')})
PY
tail -3 gpurun_out/kernels_only.log | cut -c1-300
