#!/bin/bash
# A/B of tuning builds of libsmw.so on the cfg2 bench (device-resident leg only), interleaved twice; usage: gpu_r01_ab.sh <variant> ...
mkdir -p gpurun_out
python bench.py --no-cpu --no-e2e --steps 1 --warmup 1 > /dev/null 2>&1
for rep in 1 2; do for v in "$@"; do
  if [ $v = default ]; then unset SMW_LIB; else export SMW_LIB=$PWD/stark_mlwe_b200/libsmw_$v.so; fi
  python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', round(d['value'] / 1e6, 2), 'M elems/s', round(d['ms_per_step'], 1), 'ms', {k: round(x, 1) for k, x in d['roofline']['step_share'].items()})" | tee -a gpurun_out/ab_variants.txt
done; done
