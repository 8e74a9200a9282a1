timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
timeout 600 python bench.py > gpurun_out/bench_c33_1gpu.json 2> gpurun_out/bench_c33_1gpu.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c33_1gpu.json').read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(v['ms'],2) for k,v in d['stages'].items()})"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 > gpurun_out/bench_c33_2gpu.json 2> gpurun_out/bench_c33_2gpu.err
python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_c33_2gpu.json') if l.startswith('{')][0]); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(v['ms'],2) for k,v in d['stages'].items()})"
tail -2 gpurun_out/bench_c33_2gpu.err
