#!/bin/bash
# r02-z: GEMM chain kernel: first contact.  Own timeouts everywhere: a scheduling bug in the chain shows as a trap or a hang.
set -u
O=gpurun_out/r2z
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_encoder.py -x -q -k "chain_schedule" > $O/pytest_chain.log 2>&1; echo "chain test exit $?"; tail -15 $O/pytest_chain.log
timeout 600 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 $O/pytest_gpu.log
timeout 300 python bench.py --no-cpu-baseline --no-extras > $O/bench_chain.json 2> $O/bench_chain.err; echo "bench exit $?"; tail -3 $O/bench_chain.err
FA_CHAIN=0 timeout 300 python bench.py --no-cpu-baseline --no-extras > $O/bench_nochain.json 2> $O/bench_nochain.err; echo "bench nochain exit $?"
for wl in cfg2 cfg3; do
timeout 300 python bench.py --no-cpu-baseline --no-extras --workload $wl > $O/bench_chain_$wl.json 2>> $O/bench_chain.err
FA_CHAIN=0 timeout 300 python bench.py --no-cpu-baseline --no-extras --workload $wl > $O/bench_nochain_$wl.json 2>> $O/bench_nochain.err
done
ls -la $O
