#!/bin/bash
# r02-ab: chain with deferred publication (+ L2 prefetch of the old residual rows for the f32 tiles): check + timings.
set -u
O=gpurun_out/r2ab
mkdir -p $O
run() { name=$1; shift; timeout 200 env "$@" python profiles/scripts/chain_check.py > $O/check_$name.log 2>&1; echo "$name exit $?"; grep -E "rows|OK|MISMATCH|rror" $O/check_$name.log | cut -c1-200 | tail -11; }
run default FA_X=0
run nodefer FA_CHAIN_DEFER=0
run dist48 FA_CHAIN_DIST=48
run dist16 FA_CHAIN_DIST=16
timeout 200 python profiles/scripts/with_lib.py nopf profiles/scripts/chain_check.py > $O/check_nopf.log 2>&1; echo "nopf exit $?"; grep -E "rows|OK|MISMATCH|rror" $O/check_nopf.log | cut -c1-200 | tail -11
timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_chain.json 2> $O/bench_chain.err; echo "bench exit $?"; tail -3 $O/bench_chain.err
python -c "
import json; d=json.load(open('$O/bench_chain.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['kernel_ms_per_step'], d['clocks'])"
