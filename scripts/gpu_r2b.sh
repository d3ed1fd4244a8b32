#!/bin/bash
# round 2, call b: the new GEMM (tests + A/B against the library), the tests that failed in call a, noise / loss roofline
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
export YAT_REPORT_DIR=$PWD/gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py -q -x > gpurun_out/t_gemm.log 2>&1; echo "gemm pytest rc=$?"
tail -n 25 gpurun_out/t_gemm.log | grep -v "^frame\|CUDAEvent"
timeout 900 python -m pytest tests/test_parity_k2_gpu.py tests/test_optim_gpu.py tests/test_noise_gpu.py tests/test_graph_capture_gpu.py -q > gpurun_out/t_b.log 2>&1; echo "pytest b rc=$?"
tail -n 12 gpurun_out/t_b.log | grep -v "^frame\|CUDAEvent"
timeout 600 python bench.py --kernels-only > gpurun_out/kernels_only.log 2>&1; echo "kernels-only rc=$?"
grep -E "gemm|noise|loss" gpurun_out/kernels_only.log | cut -c1-400
