#!/bin/bash
# r02-ao: evidence of the final state: ncu launch lists of one cfg5 and one cfg2 step, one `ncu --set full` capture of every hot
# kernel stand-alone at cfg5 size (roofline.traffic), kernel probe.  Usage: gpurun --timeout 1500 -- 'bash profiles/scripts/campaign_r2ao.sh'
set -u
O=gpurun_out/r2ao
mkdir -p $O
python profiles/scripts/probe_kernels.py all 5 > $O/probe_kernels.txt 2>&1; echo "probe exit $?"; cat $O/probe_kernels.txt
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
ncu -i $O/prof_kernels.ncu-rep --page source --csv > $O/ncu_source.csv 2> $O/ncu_source.err
gzip -9 $O/ncu_source.csv $O/launches_cfg5.csv $O/launches_cfg2.csv; rm -f $O/prof_kernels.ncu-rep
python profiles/scripts/ncu_summary.py $O/ncu_raw.csv > $O/ncu_kernels.md; cat $O/ncu_kernels.md
ls -la $O
