#!/bin/bash
set -u
O=gpurun_out/r2aw
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "attention" 2>&1 | tail -3
python profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 long > $O/attn.txt 2>&1; cat $O/attn.txt
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py 2>&1 | tail -2
timeout 200 python bench.py --no-cpu-baseline > $O/bench.json 2> $O/bench.err; python -c "
import json; d=json.load(open('$O/bench.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['kernel_ms_per_step']); print({k:(v.get('ms_per_step'), v.get('e2e_ms')) for k,v in d['workloads'].items()})"
