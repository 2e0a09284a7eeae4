#!/bin/bash
# round-1 GPU pass AH (8 GPUs): column-sharded cfg2 bench at N=8, cfg5-shaped run (one 2^26 column: four-step NTT over 8 ranks with one
# all-to-all, then the range-sharded FRI commit phase), range-sharded FRI of 2^22 at 8 ranks
set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/ah_bench_n$N.json 2> gpurun_out/ah_bench_n$N.err; echo "bench rc=$?"; tail -c 1200 gpurun_out/ah_bench_n$N.json; tail -n 3 gpurun_out/ah_bench_n$N.err
NTT_LOG_N=26 timeout 600 $TR --master-port 29533 tools/ntt_fri_dist_bench.py > gpurun_out/ah_dist_n${N}_2p26.json 2> gpurun_out/ah_dist_n${N}_2p26.err; echo "dist rc=$?"; cat gpurun_out/ah_dist_n${N}_2p26.json; tail -n 3 gpurun_out/ah_dist_n${N}_2p26.err
timeout 600 $TR --master-port 29535 tools/fri_shard_bench.py > gpurun_out/ah_shard_n$N.json 2> gpurun_out/ah_shard_n$N.err; echo "shard rc=$?"; cat gpurun_out/ah_shard_n$N.json; tail -n 3 gpurun_out/ah_shard_n$N.err
