run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 tools/ring_cohort_check.py --samples 12500 --bins 1000000 --k 200 --check 3 2>&1 | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['stages_rank0_ms'], d['gram_tflops_rank0_full_equiv'], d['exact_match'])
"; }
echo "== dynamic"; run 29601
echo "== dynamic"; run 29602
echo "== static all SMs"; GRIDB200_GRAM_SCHED=static run 29603
echo "== static comm16"; GRIDB200_RING_COMM_SMS=16 run 29604
echo "== dynamic nopace"; GRIDB200_GRAM_NOPACE=1 run 29605
echo "== gram_bench vmax200 dyn/static"
timeout 120 python tools/gram_bench.py 3202 2899776 200 5
GRIDB200_GRAM_SCHED=static timeout 120 python tools/gram_bench.py 3202 2899776 200 5
GRIDB200_GRAM_NOSYM=1 timeout 120 python tools/gram_bench.py 6250 1000000 200 3
GRIDB200_GRAM_NOSYM=1 GRIDB200_GRAM_SCHED=static timeout 120 python tools/gram_bench.py 6250 1000000 200 3
