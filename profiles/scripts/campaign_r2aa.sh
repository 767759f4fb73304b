#!/bin/bash
# r02-aa: where the chain loses its time: publication variants (immediate + fence, immediate, deferred), plain vs cooperative launch,
# PDL on top, dependency distance.  chain_check.py prints per-call times of a 4+2-block tower (three chained layers).
set -u
O=gpurun_out/r2aa
mkdir -p $O
run() { name=$1; shift; timeout 200 env "$@" python profiles/scripts/chain_check.py > $O/check_$name.log 2>&1; echo "$name exit $?"; grep -E "rows|OK|MISMATCH|rror" $O/check_$name.log | cut -c1-200 | tail -11; }
run deferred FA_X=0
run deferred_nocoop FA_CHAIN_COOP=0
run deferred_pdl FA_CHAIN_PDL=1
run deferred_nocoop_pdl FA_CHAIN_COOP=0 FA_CHAIN_PDL=1
run deferred_dist48 FA_CHAIN_DIST=48
run deferred_dist12 FA_CHAIN_DIST=12
for v in pub0 pub1; do
timeout 200 python profiles/scripts/with_lib.py $v profiles/scripts/chain_check.py > $O/check_$v.log 2>&1; echo "$v exit $?"; grep -E "rows|OK|MISMATCH|rror" $O/check_$v.log | cut -c1-200 | tail -11
done
