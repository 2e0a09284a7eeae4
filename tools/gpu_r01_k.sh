#!/bin/bash
# round-1 GPU pass K (2 GPUs): parity tests incl. range-sharded FRI, then cfg3 sharded over 1 and 2 ranks under torchrun (NCCL cap gather)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/k_pytest.log
tail -n 15 gpurun_out/k_pytest.log
timeout 600 python tools/fri_shard_bench.py > gpurun_out/k_fri_n1.json 2> gpurun_out/k_fri_n1.err; echo "fri1 rc=$?"; cat gpurun_out/k_fri_n1.json; tail -n 5 gpurun_out/k_fri_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 tools/fri_shard_bench.py > gpurun_out/k_fri_n2.json 2> gpurun_out/k_fri_n2.err; echo "fri2 rc=$?"; cat gpurun_out/k_fri_n2.json; tail -n 8 gpurun_out/k_fri_n2.err
