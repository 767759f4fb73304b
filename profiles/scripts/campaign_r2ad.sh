#!/bin/bash
set -u
O=gpurun_out/r2ad
mkdir -p $O
timeout 240 python profiles/scripts/chain_check.py > $O/chain_check.log 2>&1; rc=$?; echo "chain check exit $rc"; grep -E "rows|OK|MISMATCH|rror" $O/chain_check.log | cut -c1-200 | tail -12
[ $rc -ne 0 ] && exit 0
timeout 200 python profiles/scripts/with_lib.py ctrace profiles/scripts/chain_trace.py 360 0 61 > $O/trace_cfg5.txt 2>&1; echo "trace exit $?"; grep -E "===|out  :|ffn2 :|qkv  :|ffn1 :|epi w|mma:|producer:" $O/trace_cfg5.txt
timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_chain.json 2> $O/bench_chain.err; echo "bench exit $?"; tail -3 $O/bench_chain.err
python -c "
import json; d=json.load(open('$O/bench_chain.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['kernel_ms_per_step'], d['clocks'])"
