mkdir -p gpurun_out/r2bm
timeout 200 python -m pytest tests/test_gpu_encoder.py -x -q -k "int16 or overlapped" 2>&1 | tail -5
for k in 2 3 4 6 2; do
timeout 200 python bench.py --no-extras --no-cpu-baseline --e2e-pieces $k > gpurun_out/r2bm/bench_pieces$k.json 2> gpurun_out/r2bm/bench_pieces$k.err; echo "pieces $k exit $?"
python -c "
import json
d=json.load(open('gpurun_out/r2bm/bench_pieces$k.json'))
print('pieces $k dev', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['ms_per_step'],2), 'e2e16', round(d['e2e_pcm16']['ms_per_step'],2))"
done
