#!/bin/bash
set -u
O=gpurun_out/r2av
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "attention" 2>&1 | tail -3
python profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 long > $O/attn.txt 2>&1; cat $O/attn.txt
python profiles/scripts/with_lib.py psmem profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 long > $O/attn_psmem.txt 2>&1; cat $O/attn_psmem.txt
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py 2>&1 | tail -2
