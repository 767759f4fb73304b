#!/bin/bash
set -u
O=gpurun_out/r2am
mkdir -p $O
for wl in cfg2 cfg3 cfg5; do
for v in default slab1; do
if [ $v = default ]; then cmd="python bench.py"; else cmd="python profiles/scripts/with_lib.py slab1 bench.py"; fi
timeout 200 $cmd --no-cpu-baseline --no-extras --workload $wl > $O/bench_${v}_$wl.json 2> $O/err.txt; python -c "
import json; d=json.load(open('$O/bench_${v}_$wl.json')); print('$wl $v', round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), d['kernel_ms_per_step']['gemm'], d['clocks']['sm_mhz'])"
done; done
