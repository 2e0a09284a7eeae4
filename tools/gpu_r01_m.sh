#!/bin/bash
# round-1 GPU pass M: per-layer trees of the FRI commit phase on side streams: parity tests, cfg3 timings (unsharded + sharded world 1)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/m_pytest.log
tail -n 6 gpurun_out/m_pytest.log
timeout 900 python tools/fri_bench.py > gpurun_out/m_fri.json 2> gpurun_out/m_fri.err; echo "fri rc=$?"; cat gpurun_out/m_fri.json; tail -n 5 gpurun_out/m_fri.err
timeout 600 python tools/fri_shard_bench.py > gpurun_out/m_fri_n1.json 2> gpurun_out/m_fri_n1.err; echo "fri1 rc=$?"; cat gpurun_out/m_fri_n1.json; tail -n 5 gpurun_out/m_fri_n1.err
