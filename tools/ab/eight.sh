timeout 600 python -m pytest tests/test_gpu_rows.py -x -q -m gpu -k "8 or reduce_scatter" 2>&1 | tail -3
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus 8 > gpurun_out/bench_c50_8gpu.json 2> gpurun_out/bench_c50_8gpu.err
python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_c50_8gpu.json') if l.startswith('{')][0]); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(v['ms'],2) for k,v in d['stages'].items()})"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29812 bench.py --gpus 4 > gpurun_out/bench_c50_4gpu.json 2> gpurun_out/bench_c50_4gpu.err
python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_c50_4gpu.json') if l.startswith('{')][0]); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:round(v['ms'],2) for k,v in d['stages'].items()})"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29813 tools/ring_cohort_check.py --samples 50000 --bins 1500000 --k 200 --check 4 2>&1 | grep "^{" | tee gpurun_out/ring_c50_8gpu.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['stages_rank0_ms'], d['gram_tflops_rank0_full_equiv'], d['exact_match'])"
