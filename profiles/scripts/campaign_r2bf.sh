#!/bin/bash
# r02-bf: evidence of HEAD after the frontend rework, the f4 route and the allocation-free entry point: full GPU suite, smoke,
# default bench, reference arm, ncu launch list of one cfg5 step, `ncu --set full` of the two kernels that changed since r02-ao
# (frontend, attention + FSMN) alone at cfg5 size.
# Usage: gpurun --timeout 1260 -- 'bash profiles/scripts/campaign_r2bf.sh'
set -u
O=gpurun_out/r2bf
mkdir -p $O
timeout 840 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu.log
timeout 240 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
timeout 420 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
timeout 240 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref exit $?"
cp gpurun_out/cfg4_vllm_greedy.json $O/ 2>/dev/null
timeout 120 python profiles/scripts/probe_kernels.py all 5 > $O/probe_kernels.txt 2>&1; echo "probe exit $?"
timeout 120 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload cfg5 --graph-rows 0 > $O/plain_cfg5.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm|repack|splice|chain' \
    -c 3000 --csv --log-file $O/launches_cfg5.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload cfg5 --graph-rows 0 > $O/ncu_cfg5.log 2>&1
echo "ncu cfg5 exit $?"
for k in frontend attnfsmn; do
timeout 60 python profiles/scripts/probe_kernels.py $k 1 > $O/plain_probe_$k.log 2>&1 &&
timeout 240 ncu --set full --clock-control none --import-source on -k regex:'fbank|attention' -c 8 \
    -o $O/prof_$k python profiles/scripts/probe_kernels.py $k 1 > $O/ncu_full_$k.log 2>&1
echo "ncu full $k exit $?"
ncu -i $O/prof_$k.ncu-rep --page raw --csv > $O/ncu_raw_$k.csv 2> $O/ncu_raw_$k.err
done
ncu -i $O/prof_frontend.ncu-rep --page source --csv > $O/ncu_source_frontend.csv 2>/dev/null
gzip -9 $O/ncu_source_frontend.csv
rm -f $O/prof_*.ncu-rep
gzip -9 $O/launches_cfg5.csv
for k in frontend attnfsmn; do python profiles/scripts/ncu_summary.py $O/ncu_raw_$k.csv > $O/ncu_kernels_$k.md; cat $O/ncu_kernels_$k.md; done
ls -la $O
