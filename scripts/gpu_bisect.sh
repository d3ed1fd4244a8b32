#!/bin/bash
mkdir -p gpurun_out
for args in "1 256 4" "2 924 5" "16 1024 70"; do
  timeout 120 python scripts/diag_bwd.py $args > gpurun_out/bwd_diag.log 2>&1
  echo "$args rc=$? :: $(grep -v '^frame\|CUDAEvent\|^$' gpurun_out/bwd_diag.log | grep -i 'error\|rel\|done' | head -6 | cut -c1-60 | tr '\n' '|')"
done
bash scripts/gpu_second.sh
