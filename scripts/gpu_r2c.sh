#!/bin/bash
# round 2, call c: GEMM tile/ring sweep against the library on the block's projection shapes
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
timeout 600 python scripts/gemm_ab.py > gpurun_out/gemm_ab.log 2>&1; echo "gemm_ab rc=$?"
cat gpurun_out/gemm_ab.log | cut -c1-1500
