#!/bin/bash
# round-1 GPU pass H: parity tests, default bench (chunked/double-buffered e2e), reference arm, FRI commit-phase timing, launch list
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/h_pytest.log
tail -n 8 gpurun_out/h_pytest.log
timeout 900 python bench.py > gpurun_out/h_bench.json 2> gpurun_out/h_bench.err; echo "bench rc=$?"
tail -c 2600 gpurun_out/h_bench.json; tail -n 5 gpurun_out/h_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/h_bench_ref.json 2> gpurun_out/h_bench_ref.err; echo "ref rc=$?"
cat gpurun_out/h_bench_ref.json
timeout 900 python tools/fri_bench.py > gpurun_out/h_fri.json 2> gpurun_out/h_fri.err; echo "fri rc=$?"; cat gpurun_out/h_fri.json; tail -n 5 gpurun_out/h_fri.err
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/h_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/h_launches.csv $CMD > gpurun_out/h_ncu_l.log 2>&1
tail -n 3 gpurun_out/h_ncu_l.log
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/h_merkle_level -f $CMD > gpurun_out/h_ncu_f.log 2>&1
tail -n 3 gpurun_out/h_ncu_f.log
ncu --set full --clock-control none --import-source on -k regex:k_ntt -s 2 -c 3 -o gpurun_out/h_ntt -f $CMD > gpurun_out/h_ncu_n.log 2>&1
tail -n 3 gpurun_out/h_ncu_n.log
