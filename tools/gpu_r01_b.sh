#!/bin/bash
# round-1 GPU pass B: integer-pipe micro-benchmark (kept this time), parity tests, default bench, reference arm,
# then ncu: launch list + one full capture (with source) of the Merkle-level kernel on a reduced step
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/b_smi.txt
./tools/imad_peak2 > gpurun_out/b_imad_peak2.jsonl 2>&1; ./tools/imad_peak2 1048576 > gpurun_out/b_imad_peak2_long.jsonl 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
tail -5 gpurun_out/b_pytest.log
python bench.py > gpurun_out/b_bench.json 2> gpurun_out/b_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/b_bench.json
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/b_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/b_launches.csv $CMD > gpurun_out/b_ncu_l.log 2>&1
$CMD > gpurun_out/b_ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 6 -c 1 -o gpurun_out/b_merkle_level -f $CMD > gpurun_out/b_ncu_f.log 2>&1
tail -3 gpurun_out/b_ncu_l.log gpurun_out/b_ncu_f.log
