#!/bin/bash
# same-box A/B of the training step: attn2 on cuDNN (default) against attn2 on the own forward + backward kernels
mkdir -p gpurun_out
python -m yat_b200.build > gpurun_out/build.log 2>&1
F="--steps 8 --warmup 3 --no-other-workloads --no-account --no-cpu-baseline"
P='import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d["ms_per_step"],2), round(d["value"],2), round(d["e2e"]["value"],2), d["clocks"]["sm_mhz"], round(d["final_loss"],4))'
for rep in 1 2; do
echo "== library"; YAT_XATTN_OWN_BACKWARD=0 timeout 600 python bench.py $F 2>/dev/null | python -c "$P"
echo "== own";     YAT_XATTN_OWN_BACKWARD=1 timeout 600 python bench.py $F 2>/dev/null | python -c "$P"
done
