#!/bin/bash
# r01-m: single-pass attention kernel for utterances of <= 192 keys, A/B against the block-loop kernel (FA_ATTN_SHORT=0).
set -u
O=gpurun_out/m
mkdir -p $O
FA_SKIP_VLLM=1 timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 $O/pytest_gpu.log
for cfg in 0 1; do echo "== FA_ATTN_SHORT=$cfg"; FA_ATTN_SHORT=$cfg timeout 300 python profiles/scripts/probe_kernels.py attn 7 2>&1 | tail -3; done | tee $O/probe_attn.txt
for cfg in 0 1 0 1; do
  for wl in cfg5 cfg2; do
    FA_ATTN_SHORT=$cfg timeout 300 python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu-baseline > $O/bench_${wl}_short${cfg}.json 2>&1
    echo "short=$cfg $wl: $(grep -o '"ms_per_step": [0-9.]*' $O/bench_${wl}_short${cfg}.json | head -2 | tr '\n' ' ') $(grep -o '"attention": [0-9.]*' $O/bench_${wl}_short${cfg}.json | head -1)"
  done
done | tee $O/bench_summary.txt
