timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for i in 1 2; do
timeout 600 python bench.py --no-e2e --no-cpu > gpurun_out/bench_c40.json 2> gpurun_out/bench_c40.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c40.json').read()); print(d['value'], d['ms_per_step'], {k:round(v['ms'],2) for k,v in d['stages'].items()})"
done
