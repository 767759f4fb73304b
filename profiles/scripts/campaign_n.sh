#!/bin/bash
# r01-n: end-of-round state exactly as the driver runs it: full GPU suite (vLLM cfg4 included), smoke(), default bench, reference
# arm, then the ncu launch list of the same bench command.  Usage: gpurun --timeout 1800 -- 'bash profiles/scripts/campaign_n.sh'
set -u
O=gpurun_out/n
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu.log
nvidia-smi --query-compute-apps=pid,used_memory --format=csv > $O/after_pytest_apps.txt 2>&1; cat $O/after_pytest_apps.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "ref exit $?"
python bench.py --steps 5 --warmup 3 --workload cfg3 --no-cpu-baseline > $O/bench_cfg3.json 2>&1
python bench.py --steps 5 --warmup 3 --workload cfg2 --no-cpu-baseline > $O/bench_cfg2.json 2>&1
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm|repack|splice' \
    -c 6000 --csv --log-file $O/launches_cfg5.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
echo "ncu launches exit $?"
python - <<'PY'
import csv, re
rows=[r for r in csv.reader(open('gpurun_out/n/launches_cfg5.csv')) if r and not r[0].startswith('==')]
hdr=rows[0]; ix={h:i for i,h in enumerate(hdr)}
data=[r for r in rows[1:] if len(r)==len(hdr) and r[ix['Metric Name']]=='gpu__time_duration.sum']
names=[r[ix['Kernel Name']] for r in data]
fb=[i for i,n in enumerate(names) if 'fbank' in n]
seg=[(a,b) for a,b in zip(fb,fb[1:]) if b-a==442]
a,b=seg[0] if seg else (fb[0], fb[0]+442)
short=lambda n: re.sub(r'\(.*','',n).replace('void ','')
with open('gpurun_out/n/launches_cfg5_step.csv','w') as f:
    w=csv.writer(f); w.writerow(['idx','kernel','block','grid','gpu__time_duration.sum','unit'])
    for i,r in enumerate(data[a:b]):
        w.writerow([i,short(r[ix['Kernel Name']]),r[ix['Block Size']],r[ix['Grid Size']],r[ix['Metric Value']].replace(',',''),r[ix['Metric Unit']]])
print('step launches', b-a, 'segments', [y-x for x,y in zip(fb,fb[1:])])
PY
gzip -9 $O/launches_cfg5.csv
ls -la $O
