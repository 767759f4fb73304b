#!/bin/bash
# r02-aj: state after the attention rework: GPU suite (vLLM worker included), smoke, default bench with every section, reference arm.
set -u
O=gpurun_out/r2aj
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref exit $?"
cp gpurun_out/cfg4_vllm_greedy.json $O/ 2>/dev/null
ls -la $O
