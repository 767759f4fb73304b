#!/bin/bash
# r02-bo / bs: final state of round 2: full GPU suite, smoke, default bench, reference arm, ncu launch list of one cfg5 step.
# Usage: gpurun --timeout 1000 -- 'bash profiles/scripts/campaign_r2bo.sh'
set -u
O=gpurun_out/${RUN:-r2bo}
mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu.log
timeout 240 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
timeout 420 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
timeout 240 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref exit $?"
cp gpurun_out/cfg4_vllm_greedy.json $O/ 2>/dev/null
timeout 120 python profiles/scripts/probe_kernels.py all 5 > $O/probe_kernels.txt 2>&1; echo "probe exit $?"
timeout 120 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload cfg5 --graph-rows 0 > $O/plain_cfg5.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm|repack|splice|chain|pcm' \
    -c 3000 --csv --log-file $O/launches_cfg5.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload cfg5 --graph-rows 0 > $O/ncu_cfg5.log 2>&1
echo "ncu cfg5 exit $?"
gzip -9 $O/launches_cfg5.csv
ls -la $O
