#!/bin/bash
# ncu passes for round 1 (run under gpurun): launch list of a reduced bench step, then one full capture of the Merkle-level kernel
set -x
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
$CMD > gpurun_out/ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 6 -c 1 -o gpurun_out/r01_merkle_level -f $CMD > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_l.log gpurun_out/ncu_f.log
