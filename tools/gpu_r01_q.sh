#!/bin/bash
# ncu capture of the FP64-pipe Merkle-level kernel (level 0 of 16 columns x 2^18 leaves)
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/q_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/q_merkle_level52 -f $CMD > gpurun_out/q_ncu.log 2>&1
tail -n 3 gpurun_out/q_ncu.log
