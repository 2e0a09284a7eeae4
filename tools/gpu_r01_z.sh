#!/bin/bash
# round-1 GPU pass Z: t = 9 register-budget variants on cfg3, interleaved twice so that box warm-up does not favour any
set -x
mkdir -p gpurun_out
timeout 300 python tools/fri_bench.py > /dev/null 2>&1
for v in t9_160 default t9_192 t9_128 t9_224 default t9_160 t9_128 t9_192 t9_224; do
  if [ $v = default ]; then unset SMW_LIB; else export SMW_LIB=$PWD/stark_mlwe_b200/libsmw_$v.so; fi
  timeout 300 python tools/fri_bench.py 2> /dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$v', {k: round(x['ms'], 1) for k, x in d.items()})" | tee -a gpurun_out/z_t9_variants.txt
done
