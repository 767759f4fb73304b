#!/bin/bash
# r02-bg: frontend with the 12-operation power split, the pruned zero-padding butterflies and the batched mel loads:
# parity tests, smoke, A/B of the mel batching (variant melseq = one group at a time), ncu --set full + source export.
# Usage: gpurun --timeout 600 -- 'bash profiles/scripts/campaign_r2bg.sh'
set -u
O=gpurun_out/r2bg
mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_frontend.py -x -q > $O/pytest_frontend.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_frontend.log
timeout 240 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
for i in 1 2; do
timeout 60 python profiles/scripts/probe_kernels.py frontend 10 >> $O/probe_frontend.txt 2>&1; echo "probe exit $?"
true
done
cat $O/probe_frontend.txt
timeout 240 ncu --set full --clock-control none --import-source on -k regex:'fbank' -c 2 \
    -o $O/prof_frontend python profiles/scripts/probe_kernels.py frontend 1 > $O/ncu_full_frontend.log 2>&1
echo "ncu full exit $?"
ncu -i $O/prof_frontend.ncu-rep --page raw --csv > $O/ncu_raw_frontend.csv 2> /dev/null
ncu -i $O/prof_frontend.ncu-rep --page source --csv > $O/ncu_source_frontend.csv 2>/dev/null
gzip -9 $O/ncu_source_frontend.csv
rm -f $O/prof_*.ncu-rep
python profiles/scripts/ncu_summary.py $O/ncu_raw_frontend.csv > $O/ncu_kernels_frontend.md; cat $O/ncu_kernels_frontend.md
