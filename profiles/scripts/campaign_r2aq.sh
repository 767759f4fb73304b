#!/bin/bash
# r02-aq: multi-GPU: default bench (weak `value` + `strong` section) under torchrun on N GPUs of one box.
# Usage: GPURUN_ARGS="--gpus N" gpurun_retry.sh 900 'bash profiles/scripts/campaign_r2aq.sh N'
set -u
N=$1
O=gpurun_out/r2aq
mkdir -p $O
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --no-cpu-baseline > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N exit $?"
python -c "
import json; d=json.load(open('$O/bench_n$N.json')); print(d['n_gpus'], round(d['value']), round(d['ms_per_step'],2), round(d['e2e']['value']), json.dumps(d.get('strong')))"
