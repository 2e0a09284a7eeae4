#!/bin/bash
# round-1 GPU pass AC: parity suite on the re-scheduled interpreter (branch-free fused block, two terms per block), bench, ncu capture
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/ac_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/ac_pytest.log
timeout 900 python bench.py > gpurun_out/ac_bench.json 2> gpurun_out/ac_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/ac_bench.json
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/ac_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/ac_merkle_level52 -f $CMD > gpurun_out/ac_ncu_f.log 2>&1
tail -n 2 gpurun_out/ac_ncu_f.log
