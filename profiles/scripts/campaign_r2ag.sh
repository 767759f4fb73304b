#!/bin/bash
set -u
O=gpurun_out/r2ax
mkdir -p $O
for rep in 1 2; do for v in 1 0; do
FA_MEM_INTO_X=$v timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_intox${v}_$rep.json 2> $O/err.txt; python -c "
import json; d=json.load(open('$O/bench_intox${v}_$rep.json')); print('into_x=$v', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), d['kernel_ms_per_step']['gemm'], d['kernel_ms_per_step']['attention'])"
done; done
for v in 1 0; do FA_MEM_INTO_X=$v timeout 200 python bench.py --no-cpu-baseline --no-extras --workload cfg2 > $O/b2_$v.json 2>> $O/err.txt; FA_MEM_INTO_X=$v timeout 200 python bench.py --no-cpu-baseline --no-extras --workload cfg3 > $O/b3_$v.json 2>> $O/err.txt; python -c "
import json; print('into_x=$v cfg2', round(json.load(open('$O/b2_$v.json'))['ms_per_step'],3), 'cfg3', round(json.load(open('$O/b3_$v.json'))['ms_per_step'],3))"; done
