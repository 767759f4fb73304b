#!/bin/bash
set -u
O=gpurun_out/r2ba
mkdir -p $O
for v in default pf0 pf32; do
if [ $v = default ]; then cmd="python"; else cmd="python profiles/scripts/with_lib.py $v"; fi
echo "== $v"; $cmd profiles/scripts/attn_bench.py 20 cfg5 cfg3 2>&1 | tail -2
done
