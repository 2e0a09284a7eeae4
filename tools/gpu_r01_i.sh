#!/bin/bash
# round-1 GPU pass I: software-pipelined interpreter (v3): parity tests, default bench, ncu of the level-0 Merkle launch
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest.log
tail -n 15 gpurun_out/i_pytest.log
timeout 900 python bench.py --no-cpu > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?"
tail -c 2700 gpurun_out/i_bench.json; tail -n 5 gpurun_out/i_bench.err
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/i_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/i_merkle_level -f $CMD > gpurun_out/i_ncu_f.log 2>&1
tail -n 3 gpurun_out/i_ncu_f.log
