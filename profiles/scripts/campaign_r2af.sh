#!/bin/bash
set -u
O=gpurun_out/r2af
mkdir -p $O
python profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 > $O/attn_fsmn2.txt 2>&1; cat $O/attn_fsmn2.txt
for n in 4 6; do python profiles/scripts/with_lib.py fsmn$n profiles/scripts/attn_bench.py 20 cfg5 cfg2 cfg3 > $O/attn_fsmn$n.txt 2>&1; cat $O/attn_fsmn$n.txt; done
for n in 4 6; do FA_CHAIN=0 timeout 200 python profiles/scripts/with_lib.py fsmn$n bench.py --no-cpu-baseline --no-extras > $O/bench_fsmn$n.json 2> $O/bench_fsmn$n.err; python -c "
import json; d=json.load(open('$O/bench_fsmn$n.json')); print('fsmn$n', d['ms_per_step'], d['e2e']['ms_per_step'], d['kernel_ms_per_step'])"; done
