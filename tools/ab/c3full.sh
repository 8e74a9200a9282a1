run() { timeout $1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $2 tools/ring_cohort_check.py "${@:3}" 2>&1 | grep "^{" ; }
echo "== small sanity"; run 200 29821 --samples 8000 --bins 200000 --k 50 --check 3 --regen --steps 1 --warmup 1 | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['exact_match'], d['hbm_gb_rank0']); sys.exit(0 if d['exact_match'] else 1)" || exit 1
echo "== config 3 full width"; run 500 29822 --samples 50000 --bins 2900000 --k 200 --check 4 --regen --steps 2 --warmup 1 | tee gpurun_out/ring_c3_full_8gpu.json | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['samples_per_s'], d['stages_rank0_ms'], d['gram_tflops_rank0_full_equiv'], d['exact_match'], d['hbm_gb_rank0'])"
