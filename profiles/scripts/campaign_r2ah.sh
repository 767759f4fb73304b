#!/bin/bash
set -u
O=gpurun_out/r2ah
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "gemm" 2>&1 | tail -2
python profiles/scripts/probe_kernels.py gemmln > $O/probe_gemmln.txt 2>&1; cat $O/probe_gemmln.txt
for v in 1 0; do
FA_MEM_INTO_X=$v timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_intox$v.json 2> $O/bench_intox$v.err; python -c "
import json; d=json.load(open('$O/bench_intox$v.json')); print('into_x=$v', d['ms_per_step'], d['e2e']['ms_per_step'], d['kernel_ms_per_step'])"
done
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py 2>&1 | tail -2
