#!/bin/bash
# round-1 GPU pass G: interpreter-based Poseidon kernels: parity tests, default bench, ncu of the level-0 Merkle launch
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/g_pytest.log
tail -n 15 gpurun_out/g_pytest.log
timeout 900 python bench.py > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/g_bench.json; tail -n 5 gpurun_out/g_bench.err
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/g_ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/g_merkle_level -f $CMD > gpurun_out/g_ncu_f.log 2>&1
tail -n 3 gpurun_out/g_ncu_f.log
