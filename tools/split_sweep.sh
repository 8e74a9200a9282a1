#!/bin/bash
# K-split sweep of the Gram kernel at one rank's share of config 2 on 8 GPUs (--bins 362500) and at the full size.
for bins in 362500 2900000; do
  for s in 0 1 2 3 4 5 6 8 13; do
    GRIDB200_GRAM_SPLITS=$s python bench.py --bins $bins --steps 8 --warmup 2 --no-cpu --no-e2e --no-check 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('bins', $bins, 'splits', $s, 'step', round(d['ms_per_step'],3), 'gram', round(d['stages']['gram']['ms'],3), 'mhz', d['clocks']['sm_mhz'])"
  done
done
