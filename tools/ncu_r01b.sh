#!/bin/bash
set -x
./tools/imad_peak > gpurun_out/imad_peak2.jsonl 2>&1
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/ncu_plain3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/r01_merkle_level0 -f $CMD > gpurun_out/ncu_f0.log 2>&1
tail -n 3 gpurun_out/ncu_f0.log
