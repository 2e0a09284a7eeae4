#!/bin/bash
# round-1 GPU pass O: full GPU parity suite incl. the proof_bytes KAT on the GPU path
set -x
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/o_pytest.log
tail -n 25 gpurun_out/o_pytest.log
