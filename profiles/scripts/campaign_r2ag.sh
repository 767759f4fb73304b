#!/bin/bash
set -u
O=gpurun_out/r2bd
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "attention" 2>&1 | tail -2
for v in default nol2; do
if [ $v = default ]; then cmd="python"; else cmd="python profiles/scripts/with_lib.py $v"; fi
echo "== $v"; $cmd profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 long 2>&1 | tail -4
done
