#!/bin/bash
# round-1 GPU pass J (2 GPUs): parity tests, default bench at N=1, bench at N=2 under torchrun (column sharding + NCCL root all-gather)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/j_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/j_pytest.log
tail -n 5 gpurun_out/j_pytest.log
timeout 900 python bench.py --no-cpu > gpurun_out/j_bench1.json 2> gpurun_out/j_bench1.err; echo "bench1 rc=$?"
tail -c 2700 gpurun_out/j_bench1.json; tail -n 5 gpurun_out/j_bench1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/j_bench2.json 2> gpurun_out/j_bench2.err; echo "bench2 rc=$?"
tail -c 2700 gpurun_out/j_bench2.json; tail -n 8 gpurun_out/j_bench2.err
