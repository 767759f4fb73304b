#!/bin/bash
# r02-y: out-projection without the identity block (ABI 7): op/encoder/bench-path parity, default bench, GEMM probes for the
# product build and the one-slab bf16-epilogue variant (one more operand stage), ncu launch lists of a cfg5 and a cfg2 step.
# Usage: gpurun --timeout 1500 -- 'bash profiles/scripts/campaign_r2y.sh'
set -u
O=gpurun_out/r2y
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_zz_vllm_greedy.py > $O/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $O/pytest_gpu.log
python bench.py --no-cpu-baseline > $O/bench_default.json 2> $O/bench_default.err; echo "bench exit $?"
python profiles/scripts/probe_kernels.py gemm > $O/probe_gemm.txt 2>&1; python profiles/scripts/probe_kernels.py gemmln >> $O/probe_gemm.txt 2>&1
python profiles/scripts/with_lib.py slab1 profiles/scripts/probe_kernels.py gemm > $O/probe_gemm_slab1.txt 2>&1
python profiles/scripts/with_lib.py slab1 profiles/scripts/probe_kernels.py gemmln >> $O/probe_gemm_slab1.txt 2>&1
python profiles/scripts/with_lib.py slab1 bench.py --no-cpu-baseline --no-extras > $O/bench_slab1.json 2> $O/bench_slab1.err
python bench.py --no-cpu-baseline --no-extras > $O/bench_default_noextras.json 2>> $O/bench_default.err
for wl in cfg5 cfg2; do
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm|repack|splice' \
    -c 3000 --csv --log-file $O/launches_$wl.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --workload $wl --graph-rows 0 > $O/ncu_$wl.log 2>&1
echo "ncu $wl exit $?"
done
gzip -9 $O/launches_cfg5.csv $O/launches_cfg2.csv
ls -la $O
