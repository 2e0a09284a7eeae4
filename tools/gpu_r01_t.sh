#!/bin/bash
# round-1 GPU pass T: final-form bench (default run incl. cpu_baseline), reference arm, smoke, launch list + full ncu of the FP64 Merkle kernel with one-warp CTAs
set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/t_smoke.log
timeout 900 python bench.py > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err; echo "bench rc=$?"
tail -c 3500 gpurun_out/t_bench.json; tail -n 5 gpurun_out/t_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/t_bench_ref.json 2> gpurun_out/t_bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e"
$CMD > gpurun_out/t_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/t_launches.csv $CMD > gpurun_out/t_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_merkle_level -s 5 -c 1 -o gpurun_out/t_merkle_level52 -f $CMD > gpurun_out/t_ncu_f.log 2>&1
tail -n 2 gpurun_out/t_ncu_f.log
