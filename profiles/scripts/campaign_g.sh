#!/bin/bash
# r01-g: overlapped e2e schedule (two half-batches, H2D/D2H under compute), schedule sweeps, cfg4 vLLM greedy agreement.
# Usage: gpurun --timeout 1500 -- 'bash profiles/scripts/campaign_g.sh'
set -u
O=gpurun_out/g
mkdir -p $O
FA_SKIP_VLLM=1 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > $O/bench_cfg5.json 2> $O/bench_cfg5.err; echo "cfg5 exit $?"
python bench.py --steps 5 --warmup 3 --no-overlap --no-cpu-baseline > $O/bench_cfg5_nooverlap.json 2>&1; echo "cfg5 no-overlap exit $?"
python bench.py --steps 5 --warmup 3 --streams 1 --no-cpu-baseline > $O/bench_cfg5_streams1.json 2>&1; echo "cfg5 streams1 exit $?"
python bench.py --steps 5 --warmup 3 --chunk-tokens 30060 --no-cpu-baseline > $O/bench_cfg5_chunk30k.json 2>&1; echo "chunk30k exit $?"
python bench.py --steps 5 --warmup 3 --chunk-tokens 15030 --streams 1 --no-cpu-baseline > $O/bench_cfg5_chunk15k.json 2>&1; echo "chunk15k exit $?"
python bench.py --steps 5 --warmup 3 --workload cfg3 --no-cpu-baseline > $O/bench_cfg3.json 2>&1; echo "cfg3 exit $?"
python bench.py --steps 5 --warmup 3 --workload cfg2 --no-cpu-baseline > $O/bench_cfg2.json 2>&1; echo "cfg2 exit $?"
grep -ho '"ms_per_step": [0-9.]*' $O/bench_cfg5*.json | paste - - - | head
timeout 1000 python -m pytest tests/test_gpu_zz_vllm_greedy.py -m gpu -x -q -s > $O/cfg4_pytest.log 2>&1; echo "cfg4 exit $?"
tail -5 $O/cfg4_pytest.log
ls -la $O
