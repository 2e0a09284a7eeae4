#!/bin/bash
# round-1 GPU pass L (2 GPUs): parity tests incl. distributed NTT, then the cfg5-shaped NTT+FRI run at 1 and 2 ranks (NCCL all-to-all + cap gathers)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/l_pytest.log
tail -n 15 gpurun_out/l_pytest.log
timeout 600 python tools/ntt_fri_dist_bench.py > gpurun_out/l_dist_n1.json 2> gpurun_out/l_dist_n1.err; echo "dist1 rc=$?"; cat gpurun_out/l_dist_n1.json; tail -n 5 gpurun_out/l_dist_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/ntt_fri_dist_bench.py > gpurun_out/l_dist_n2.json 2> gpurun_out/l_dist_n2.err; echo "dist2 rc=$?"; cat gpurun_out/l_dist_n2.json; tail -n 12 gpurun_out/l_dist_n2.err
NTT_LOG_N=26 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29523 tools/ntt_fri_dist_bench.py > gpurun_out/l_dist_n2_2p26.json 2> gpurun_out/l_dist_n2_2p26.err; echo "dist2-26 rc=$?"; cat gpurun_out/l_dist_n2_2p26.json; tail -n 12 gpurun_out/l_dist_n2_2p26.err
