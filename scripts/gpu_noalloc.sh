#!/bin/bash
# the no-allocation entry point: alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q 2>&1 | tail -5
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -8
