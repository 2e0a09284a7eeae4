#!/bin/bash
# round-1 GPU pass W (2 GPUs): the final kernels at N=2 — column-sharded bench (both arms), range-sharded FRI, distributed NTT + FRI at 2^26
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/w_bench_n2.json 2> gpurun_out/w_bench_n2.err; echo "bench2 rc=$?"; tail -c 1500 gpurun_out/w_bench_n2.json; tail -n 4 gpurun_out/w_bench_n2.err
timeout 600 python tools/fri_shard_bench.py > gpurun_out/w_shard_n1.json 2> gpurun_out/w_shard_n1.err; echo "shard1 rc=$?"; cat gpurun_out/w_shard_n1.json
timeout 600 $TR --master-port 29513 tools/fri_shard_bench.py > gpurun_out/w_shard_n2.json 2> gpurun_out/w_shard_n2.err; echo "shard2 rc=$?"; cat gpurun_out/w_shard_n2.json; tail -n 4 gpurun_out/w_shard_n2.err
NTT_LOG_N=26 timeout 600 $TR --master-port 29515 tools/ntt_fri_dist_bench.py > gpurun_out/w_dist_n2_2p26.json 2> gpurun_out/w_dist_n2_2p26.err; echo "dist2-26 rc=$?"; cat gpurun_out/w_dist_n2_2p26.json; tail -n 4 gpurun_out/w_dist_n2_2p26.err
