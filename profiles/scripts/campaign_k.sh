#!/bin/bash
# r01-k: serpentine schedule A/B (FA_SERPENTINE=1: every other launch walks its rows last-to-first so it starts on what its
# predecessor wrote last, still in L2).  Usage: gpurun --timeout 900 -- 'bash profiles/scripts/campaign_k.sh'
set -u
O=gpurun_out/k
mkdir -p $O
FA_SERPENTINE=1 FA_SKIP_VLLM=1 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_serp.log 2>&1; echo "pytest serp exit $?"; tail -2 $O/pytest_gpu_serp.log
for serp in 0 1 0 1; do
  for wl in cfg5 cfg3 cfg2; do
    FA_SERPENTINE=$serp python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu-baseline > $O/bench_${wl}_serp${serp}.json 2>&1
    echo "serp=$serp $wl: $(grep -o '"ms_per_step": [0-9.]*' $O/bench_${wl}_serp${serp}.json | head -2 | tr '\n' ' ') $(grep -o '"sm_mhz": [0-9.]*' $O/bench_${wl}_serp${serp}.json) $(grep -o '"kernel_ms_per_step": {[^}]*}' $O/bench_${wl}_serp${serp}.json)"
  done
done | tee $O/bench_summary.txt
