#!/bin/bash
# r02-be: evidence of the state after the frontend rework and the f4 route: full GPU suite, smoke, default bench, reference arm,
# ncu launch lists of one cfg5 and one cfg2 step, one `ncu --set full` capture of every hot kernel alone at cfg5 size.
# Usage: gpurun --timeout 1500 -- 'bash profiles/scripts/campaign_r2be.sh'
set -u
O=gpurun_out/r2be
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref exit $?"
cp gpurun_out/cfg4_vllm_greedy.json $O/ 2>/dev/null
python profiles/scripts/probe_kernels.py all 5 > $O/probe_kernels.txt 2>&1; echo "probe exit $?"
for wl in cfg5 cfg2; do
python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload $wl --graph-rows 0 > $O/plain_$wl.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm|repack|splice|chain' \
    -c 3000 --csv --log-file $O/launches_$wl.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload $wl --graph-rows 0 > $O/ncu_$wl.log 2>&1
echo "ncu $wl exit $?"
done
python profiles/scripts/probe_kernels.py all 1 > $O/plain_probe.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fbank|gemm_bf16|attention|layernorm' -c 40 \
    -o $O/prof_kernels python profiles/scripts/probe_kernels.py all 1 > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
ncu -i $O/prof_kernels.ncu-rep --page raw --csv > $O/ncu_raw.csv 2> $O/ncu_raw.err
rm -f $O/prof_kernels.ncu-rep
gzip -9 $O/launches_cfg5.csv $O/launches_cfg2.csv
python profiles/scripts/ncu_summary.py $O/ncu_raw.csv > $O/ncu_kernels.md; cat $O/ncu_kernels.md
ls -la $O
