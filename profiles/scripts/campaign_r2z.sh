#!/bin/bash
# r02-z: GEMM chain kernel.  Own timeouts everywhere, and stop at the first failure: a scheduling bug in the chain shows as a
# trap (watchdog in every wait of the kernel) or a mismatch.
set -u
O=gpurun_out/r2z
mkdir -p $O
timeout 240 python profiles/scripts/chain_check.py > $O/chain_check.log 2>&1; rc=$?; echo "chain check exit $rc"; tail -14 $O/chain_check.log
[ $rc -ne 0 ] && exit 0
timeout 200 python -m pytest tests/test_gpu_encoder.py -x -q -k "chain_schedule" > $O/pytest_chain.log 2>&1; rc=$?; echo "chain test exit $rc"; tail -5 $O/pytest_chain.log
[ $rc -ne 0 ] && exit 0
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py > $O/pytest_gpu.log 2>&1; rc=$?; echo "pytest exit $rc"; tail -5 $O/pytest_gpu.log
[ $rc -ne 0 ] && exit 0
timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_chain.json 2> $O/bench_chain.err; echo "bench exit $?"; tail -3 $O/bench_chain.err
FA_CHAIN=0 timeout 200 python bench.py --no-cpu-baseline --no-extras > $O/bench_nochain.json 2> $O/bench_nochain.err; echo "bench nochain exit $?"
for wl in cfg2 cfg3; do
timeout 200 python bench.py --no-cpu-baseline --no-extras --workload $wl > $O/bench_chain_$wl.json 2>> $O/bench_chain.err
FA_CHAIN=0 timeout 200 python bench.py --no-cpu-baseline --no-extras --workload $wl > $O/bench_nochain_$wl.json 2>> $O/bench_nochain.err
done
ls -la $O
