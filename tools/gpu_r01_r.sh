#!/bin/bash
# threads-per-CTA sweep of the Merkle kernels (t = 17) on 16 columns x 2^18 rows, blow-up 4: FP64-pipe and integer-pipe multipliers
mkdir -p gpurun_out
run() { python bench.py --steps 2 --warmup 2 --log-n 18 --cols 16 --no-cpu --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['roofline']['step_share']['merkle_level'])"; }
for nt in 128 96 64 32; do echo "f52 nt=$nt $(SMW_NT52=$nt run)"; done | tee gpurun_out/r_nt_sweep2.txt
for nt in 160 96 64 32; do echo "f29 nt=$nt $(SMW_REP=29 SMW_NT29=$nt run)"; done | tee -a gpurun_out/r_nt_sweep2.txt
