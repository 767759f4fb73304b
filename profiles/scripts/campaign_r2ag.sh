#!/bin/bash
set -u
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "attention" 2>&1 | tail -2
python profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 long 2>&1 | tail -4
