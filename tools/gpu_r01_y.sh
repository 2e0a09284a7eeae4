#!/bin/bash
# round-1 GPU pass Y: t = 9 register-budget variants on cfg3 (fold-2 = pair leaves + arity-2 trees at t = 9), and a cfg4-sized slice
# (8 columns of 2^24 rows, blow-up 8 -> 2^27-leaf trees) through the bench pipeline, device-resident and host-streamed
set -x
mkdir -p gpurun_out
for v in default t9_192 t9_160; do
  if [ $v = default ]; then unset SMW_LIB; else export SMW_LIB=$PWD/stark_mlwe_b200/libsmw_$v.so; fi
  timeout 300 python tools/fri_bench.py > gpurun_out/y_fri_$v.json 2> gpurun_out/y_fri_$v.err; echo "$v rc=$?"; tail -c 600 gpurun_out/y_fri_$v.json
done
unset SMW_LIB
timeout 900 python bench.py --log-n 24 --log-blowup 3 --cols 8 --steps 1 --warmup 1 --no-cpu > gpurun_out/y_cfg4_slice.json 2> gpurun_out/y_cfg4_slice.err; echo "cfg4 rc=$?"
tail -c 2500 gpurun_out/y_cfg4_slice.json; tail -n 5 gpurun_out/y_cfg4_slice.err
