#!/bin/bash
# CTA-size defaults: parity suite, cfg3 sweep over CTA sizes (t = 9 kernels dominate fold-2), then the cfg2 bench with both multipliers
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/s_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s_pytest.log
tail -n 4 gpurun_out/s_pytest.log
fri() { python tools/fri_bench.py 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print({k:round(v['ms'],1) for k,v in d.items()})"; }
for nt in 32 64 128 256; do echo "f52 nt=$nt $(SMW_NT52=$nt fri)"; done | tee gpurun_out/s_fri_sweep.txt
for nt in 32 64 128 256; do echo "f29 nt=$nt $(SMW_REP=29 SMW_NT29=$nt fri)"; done | tee -a gpurun_out/s_fri_sweep.txt
timeout 900 python bench.py --no-cpu > gpurun_out/s_bench_f52.json 2> gpurun_out/s_bench_f52.err; echo "bench52 rc=$?"
SMW_REP=29 timeout 900 python bench.py --no-cpu > gpurun_out/s_bench_f29.json 2> gpurun_out/s_bench_f29.err; echo "bench29 rc=$?"
for f in s_bench_f52 s_bench_f29; do python -c "
import json
d=json.loads([l for l in open('gpurun_out/$f.json').read().splitlines() if l.startswith('{')][-1])
print('$f',d['value'],d['e2e']['value'],d['roofline']['step_share'],d['clocks'])"; done
