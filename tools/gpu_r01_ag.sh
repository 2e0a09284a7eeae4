#!/bin/bash
# round-1 GPU pass AG: the integer-multiplier interpreter (SMW_REP=29; also what widths 33/65/129 and the verifier use) with the
# branch-free fused block and two terms per block, A/B on the cfg2 bench, then the parity suite on the variant
mkdir -p gpurun_out
export SMW_REP=29
python bench.py --no-cpu --no-e2e --steps 1 --warmup 1 > /dev/null 2>&1
for rep in 1 2; do for v in default f29v; do
  if [ $v = default ]; then unset SMW_LIB; else export SMW_LIB=$PWD/stark_mlwe_b200/libsmw_$v.so; fi
  python bench.py --no-cpu --no-e2e --steps 2 --warmup 1 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('REP29 $v', round(d['value'] / 1e6, 2), 'M elems/s', round(d['ms_per_step'], 1), 'ms')" | tee -a gpurun_out/ag_variants.txt
done; done
unset SMW_REP
export SMW_LIB=$PWD/stark_mlwe_b200/libsmw_f29v.so
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3 | tee -a gpurun_out/ag_variants.txt
SMW_REP=29 timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3 | tee -a gpurun_out/ag_variants.txt
