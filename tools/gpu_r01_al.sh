#!/bin/bash
# round-1 GPU pass AL: last verification of HEAD (parity suite, smoke, default bench) + one ncu capture of the t = 9 pair-leaf kernel (cfg3's leaf hash)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/al_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/al_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/al_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/al_smoke.log
timeout 900 python bench.py > gpurun_out/al_bench.json 2> gpurun_out/al_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/al_bench.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_pair_leaves52 -s 2 -c 1 -o gpurun_out/al_pair_leaves52 -f python tools/fri_rep_times.py > gpurun_out/al_ncu.log 2>&1
tail -n 2 gpurun_out/al_ncu.log
