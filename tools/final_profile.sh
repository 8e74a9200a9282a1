#!/bin/bash
# Round-end evidence on ONE GPU: the GPU suite, the bench line, the ncu launch list of the bench command and one
# `--set full` capture of the step's five main kernels.  Usage: bash tools/final_profile.sh <tag>   (writes gpurun_out/<tag>_*)
tag=${1:-final}
out=gpurun_out
python -m pytest tests -x -q -m gpu > $out/${tag}_pytest_gpu.log 2>&1; tail -2 $out/${tag}_pytest_gpu.log
python bench.py --steps 10 --warmup 3 > $out/${tag}_bench_1gpu.json 2> $out/${tag}_bench_1gpu.err; echo "bench rc=$?"
prof="python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --no-check"
$prof > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file $out/${tag}_ncu_launches.csv $prof > $out/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
one="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-check"
ncu --set full --clock-control none --import-source on -k 'regex:k_gram_tc_pair|k_col_raw_u16|k_col_norm_u16|k_transform|k_topk_merge' \
  --launch-skip 5 -c 5 -f -o $out/${tag}_ncu_full $one > $out/${tag}_ncu_full.log 2>&1; echo "full capture rc=$?"
ls -la $out/${tag}_*
