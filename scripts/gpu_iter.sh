#!/bin/bash
# iteration call: given test files, kernel table, short bench with in-situ kernel times, ncu of glumb + optim
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1 || { cat gpurun_out/build.log; exit 1; }
timeout 1200 python -m pytest ${@:-tests} -q -m gpu -x 2>&1 | tail -12 | grep -v "^frame\|CUDAEvent"
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.log 2> gpurun_out/bench_quick.err; echo "bench rc=$?"
python -c "
import json
for l in open('gpurun_out/bench_quick.log'):
    if l.startswith('{'):
        d = json.loads(l); print('value', round(d['value'], 2), 'e2e', round(d['e2e']['value'], 2), 'ms/step', round(d['ms_per_step'], 1), 'launches', d['gpu_launches'], 'ckpt', d['config'].get('gradient_checkpointing'), 'loss', d['final_loss'])
        for k, v in d['kernels']['in_situ'].items(): print('  ', k, round(v['us'], 1), 'us x', v['launches_per_step'], 'share', round(v['share_of_step'], 4), 'frac', round(v.get('frac', 0), 3))
"
grep -v "^frame\|CUDAEvent" gpurun_out/bench_quick.err | tail -n 8
timeout 300 python scripts/prof_glumb.py > gpurun_out/plain_glumb.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"glumb_fwd|glumb_bwd" -s 2 -c 2 \
    -o gpurun_out/prof_glumb -f python scripts/prof_glumb.py > gpurun_out/ncu_glumb.log 2>&1
echo "glumb capture rc=$?"; tail -n 2 gpurun_out/ncu_glumb.log
