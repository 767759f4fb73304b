#!/bin/bash
set -u
O=gpurun_out/r2ap
mkdir -p $O
for wl in cfg2 cfg3; do
for v in default fx96 fx160; do
if [ $v = default ]; then cmd="python bench.py"; else cmd="python profiles/scripts/with_lib.py $v bench.py"; fi
timeout 200 $cmd --no-cpu-baseline --no-extras --workload $wl > $O/bench_${v}_$wl.json 2> $O/err.txt; python -c "
import json; d=json.load(open('$O/bench_${v}_$wl.json')); print('$wl $v', round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), d['kernel_ms_per_step']['gemm'])"
done; done
