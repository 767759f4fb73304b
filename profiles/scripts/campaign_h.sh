#!/bin/bash
# r01-h: shared-memory staged FSMN kernel A/B (register-window kernel vs staged, 32 / 64 rows per CTA), overlapped e2e default.
# Usage: gpurun --timeout 1200 -- 'bash profiles/scripts/campaign_h.sh'
set -u
O=gpurun_out/h
mkdir -p $O
FA_SKIP_VLLM=1 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
FA_FSMN_ROWS=64 FA_SKIP_VLLM=1 python -m pytest tests/test_gpu_ops.py tests/test_gpu_encoder.py -m gpu -x -q -k "fsmn or oracle" > $O/pytest_rows64.log 2>&1; echo "pytest rows64 exit $?"
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k fsmn > $O/sanitizer_fsmn.log 2>&1; echo "sanitizer exit $?"; tail -3 $O/sanitizer_fsmn.log
for cfg in "FA_FSMN_STAGED=0" "FA_FSMN_ROWS=32" "FA_FSMN_ROWS=64"; do
  echo "== $cfg"; env $cfg python profiles/scripts/probe_kernels.py fsmn 7 2>&1 | tail -1
done | tee $O/probe_fsmn.txt
for cfg in "FA_FSMN_STAGED=0" "FA_FSMN_ROWS=32" "FA_FSMN_ROWS=64"; do
  for wl in cfg5 cfg3 cfg2; do
    env $cfg python bench.py --steps 5 --warmup 3 --workload $wl --no-cpu-baseline > $O/bench_${wl}_${cfg//=/_}.json 2>&1
    echo "$cfg $wl: $(grep -o '"ms_per_step": [0-9.]*' $O/bench_${wl}_${cfg//=/_}.json | head -2 | tr '\n' ' ') $(grep -o '"fsmn": [0-9.]*' $O/bench_${wl}_${cfg//=/_}.json | head -1)"
  done
done | tee $O/bench_summary.txt
ls -la $O
