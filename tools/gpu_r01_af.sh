#!/bin/bash
# round-1 GPU pass AF: full parity suite on the final kernels, smoke, default bench + reference arm, launch list + full ncu capture
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/af_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/af_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/af_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/af_smoke.log
timeout 900 python bench.py > gpurun_out/af_bench.json 2> gpurun_out/af_bench.err; echo "bench rc=$?"
tail -c 3800 gpurun_out/af_bench.json; tail -n 5 gpurun_out/af_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/af_bench_ref.json 2> gpurun_out/af_bench_ref.err; echo "ref rc=$?"; tail -c 600 gpurun_out/af_bench_ref.json
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/af_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/af_launches.csv $CMD > gpurun_out/af_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/af_merkle_level52 -f $CMD > gpurun_out/af_ncu_f.log 2>&1
tail -n 2 gpurun_out/af_ncu_f.log
