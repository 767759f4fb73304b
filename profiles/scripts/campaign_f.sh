#!/bin/bash
# r01-f measurement campaign (one gpurun call, one B200): GPU parity tests, bench lines, standalone kernel probe,
# ncu launch list of one bench step and one `ncu --set full` capture of every hot kernel (for roofline.traffic).
# Usage: gpurun --timeout 1700 -- 'bash profiles/scripts/campaign_f.sh'
set -u
O=gpurun_out/f
mkdir -p $O
FA_SKIP_VLLM=1 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python bench.py --steps 5 --warmup 3 > $O/bench_cfg5.json 2> $O/bench_cfg5.err; echo "cfg5 exit $?"
python bench.py --steps 5 --warmup 3 --workload cfg3 --no-cpu-baseline > $O/bench_cfg3.json 2> $O/bench_cfg3.err; echo "cfg3 exit $?"
python bench.py --steps 5 --warmup 3 --workload cfg2 --no-cpu-baseline > $O/bench_cfg2.json 2> $O/bench_cfg2.err; echo "cfg2 exit $?"
python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref exit $?"
python profiles/scripts/probe_kernels.py all 5 > $O/probe_kernels.txt 2>&1; echo "probe exit $?"
cat $O/probe_kernels.txt
# ncu: launch list of one cfg5 step (plain run with the same arguments directly before, && chained)
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/plain_launches.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm|repack|splice' \
    -c 6000 --csv --log-file $O/launches_cfg5.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_launches.log 2>&1
echo "ncu launches exit $?"
# ncu --set full: every hot kernel once, standalone at cfg5 size
python profiles/scripts/probe_kernels.py all 1 > $O/plain_probe.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'fbank|gemm_bf16|attention|fsmn|layernorm' -c 44 \
    -o $O/prof_kernels python profiles/scripts/probe_kernels.py all 1 > $O/ncu_full.log 2>&1
echo "ncu full exit $?"
# the .ncu-rep is larger than gpurun's 64 MiB return limit: export the pages here and drop the report
ncu -i $O/prof_kernels.ncu-rep --page raw --csv > $O/ncu_raw.csv 2> $O/ncu_raw.err
ncu -i $O/prof_kernels.ncu-rep --page source --csv > $O/ncu_source.csv 2> $O/ncu_source.err
gzip -9 $O/ncu_source.csv; rm -f $O/prof_kernels.ncu-rep
# cfg4: vLLM greedy agreement (own process, hard timeout)
timeout 1000 python -m pytest tests/test_gpu_zz_vllm_greedy.py -m gpu -x -q -s > $O/cfg4_pytest.log 2>&1; echo "cfg4 exit $?"
tail -5 $O/cfg4_pytest.log
du -sh gpurun_out
ls -la $O
