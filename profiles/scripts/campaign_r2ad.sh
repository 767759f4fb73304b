#!/bin/bash
set -u
O=gpurun_out/r2ae
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_ops.py -x -q -k "attention" > $O/pytest_attn.log 2>&1; rc=$?; echo "attention ops exit $rc"; tail -3 $O/pytest_attn.log
[ $rc -ne 0 ] && exit 0
timeout 240 python profiles/scripts/chain_check.py > $O/chain_check.log 2>&1; rc=$?; echo "chain check exit $rc"; grep -E "rows|OK|MISMATCH|rror" $O/chain_check.log | cut -c1-200 | tail -12
[ $rc -ne 0 ] && exit 0
for dist in 24 48; do
FA_CHAIN_DIST=$dist timeout 200 python profiles/scripts/with_lib.py ctrace profiles/scripts/chain_trace.py 360 0 60 > $O/trace_cfg5_d$dist.txt 2>&1; echo "trace d$dist exit $?"; grep -E "===|out  :|ffn2 :|qkv  :|ffn1 :|epi w2|mma:|producer:" $O/trace_cfg5_d$dist.txt | head -9
done
FA_CHAIN_DIST=48 timeout 240 python profiles/scripts/chain_check.py 2>&1 | grep -E "rows  60120" | cut -c1-200
timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_chain.json 2> $O/bench_chain.err; echo "bench exit $?"; tail -3 $O/bench_chain.err
FA_CHAIN=0 timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_nochain.json 2> $O/bench_nochain.err; echo "bench exit $?"
python -c "
import json
for f in ('bench_chain','bench_nochain'):
    d=json.load(open('$O/'+f+'.json')); print(f, d['ms_per_step'], d['e2e']['ms_per_step'], d['kernel_ms_per_step'], d['clocks'])"
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu.log
