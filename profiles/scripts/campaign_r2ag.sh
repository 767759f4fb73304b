#!/bin/bash
set -u
O=gpurun_out/r2bb
mkdir -p $O
for v in default hint hint1k; do
if [ $v = default ]; then cmd="python"; else cmd="python profiles/scripts/with_lib.py $v"; fi
echo "== $v"; $cmd profiles/scripts/attn_bench.py 20 cfg5 cfg2 2>&1 | tail -2
timeout 200 $cmd bench.py --no-cpu-baseline --no-extras > $O/bench_$v.json 2> $O/err.txt; python -c "
import json; d=json.load(open('$O/bench_$v.json')); print('$v', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), d['kernel_ms_per_step']['gemm'], d['kernel_ms_per_step']['attention'], d['clocks']['sm_mhz'])"
timeout 200 $cmd bench.py --no-cpu-baseline --no-extras --workload cfg2 > $O/b2_$v.json 2>> $O/err.txt; python -c "
import json; print('$v cfg2', round(json.load(open('$O/b2_$v.json'))['ms_per_step'],3))"
done
