#!/bin/bash
set -u
O=gpurun_out/r2az
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "gemm or layernorm" 2>&1 | tail -2
python profiles/scripts/probe_kernels.py gemmln 2>&1 | grep "out+ln"
FA_XBS=0 python profiles/scripts/probe_kernels.py gemmln 2>&1 | grep "out+ln"
for rep in 1 2; do for v in 1 0; do
FA_XBS=$v timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_xbs${v}_$rep.json 2> $O/err.txt; python -c "
import json; d=json.load(open('$O/bench_xbs${v}_$rep.json')); print('xbs=$v', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2), d['kernel_ms_per_step']['gemm'], d['kernel_ms_per_step']['attention'])"
done; done
for v in 1 0; do FA_XBS=$v timeout 200 python bench.py --no-cpu-baseline --no-extras --workload cfg2 > $O/b2_$v.json 2>> $O/err.txt; FA_XBS=$v timeout 200 python bench.py --no-cpu-baseline --no-extras --workload cfg3 > $O/b3_$v.json 2>> $O/err.txt; python -c "
import json; print('xbs=$v cfg2', round(json.load(open('$O/b2_$v.json'))['ms_per_step'],3), 'cfg3', round(json.load(open('$O/b3_$v.json'))['ms_per_step'],3))"; done
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py 2>&1 | tail -2
