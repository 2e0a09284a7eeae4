#!/bin/bash
# round-1 GPU pass P: FP64-pipe multiplier (fr52) in the batch Poseidon kernels: parity suite, then the bench with both multipliers
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/p_pytest.log
tail -n 25 gpurun_out/p_pytest.log
timeout 900 python bench.py --no-cpu > gpurun_out/p_bench_f52.json 2> gpurun_out/p_bench_f52.err; echo "bench52 rc=$?"
tail -c 2700 gpurun_out/p_bench_f52.json; tail -n 5 gpurun_out/p_bench_f52.err
SMW_REP=29 timeout 900 python bench.py --no-cpu > gpurun_out/p_bench_f29.json 2> gpurun_out/p_bench_f29.err; echo "bench29 rc=$?"
tail -c 600 gpurun_out/p_bench_f29.json | head -c 600
