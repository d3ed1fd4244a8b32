#!/bin/bash
# experiment: does per-shape cuBLASLt algorithm selection (torch TunableOp) shorten the power-capped step?  Same-box A/B.
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
F="--steps 8 --warmup 3 --no-other-workloads --no-account --no-cpu-baseline"
echo "== baseline"; timeout 600 python bench.py $F 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])"
export PYTORCH_TUNABLEOP_ENABLED=1 PYTORCH_TUNABLEOP_TUNING=1 PYTORCH_TUNABLEOP_FILENAME=$PWD/gpurun_out/tunableop_results.csv
export PYTORCH_TUNABLEOP_MAX_TUNING_DURATION_MS=40 PYTORCH_TUNABLEOP_MAX_WARMUP_DURATION_MS=5 PYTORCH_TUNABLEOP_VERBOSE=0
echo "== tuning run"; timeout 1500 python bench.py --steps 2 --warmup 2 --no-other-workloads --no-account --no-cpu-baseline > gpurun_out/tune.log 2>&1; echo "rc=$?"; tail -c 300 gpurun_out/tune.log
ls -la gpurun_out/tunableop_results*.csv; wc -l gpurun_out/tunableop_results*.csv
export PYTORCH_TUNABLEOP_TUNING=0
echo "== tuned"; timeout 600 python bench.py $F 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])"
unset PYTORCH_TUNABLEOP_ENABLED PYTORCH_TUNABLEOP_TUNING
echo "== baseline again"; timeout 600 python bench.py $F 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['value'], d['e2e']['value'], d['clocks'])"
