#!/bin/bash
mkdir -p gpurun_out
for w in noise loss; do timeout 120 python scripts/diag_fwd.py $w > gpurun_out/diag_$w.log 2>&1; echo "$w rc=$?"; tail -2 gpurun_out/diag_$w.log; done
timeout 200 python scripts/diag_fwd.py fwd 1 4 8 9 12 16 32 > gpurun_out/diag_fwd.log 2>&1; echo "fwd rc=$?"; grep -v "^frame\|CUDAEvent" gpurun_out/diag_fwd.log | head -40
