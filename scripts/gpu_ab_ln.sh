#!/bin/bash
python -m yat_b200.build > /dev/null
for v in cols pairs cols pairs; do
  if [ $v = pairs ]; then export YAT_LN_BWD_PAIRS=1; else unset YAT_LN_BWD_PAIRS; fi
  timeout 600 python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); k = d['kernels']['in_situ']['ln_modulate_bwd']
        print('$v', 'ms/step', round(d['ms_per_step'], 2), 'ln_bwd us', round(k['us'], 1), 'clk', d['clocks']['sm_mhz'])
"
done
