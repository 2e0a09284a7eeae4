#!/bin/bash
# One parameterised runner for a GPU lease (replaces the per-lease scripts of round 1):
#   gpurun --timeout 1500 -- 'bash tools/gpu_run.sh <tag> <stage> [<stage> ...]'
# Every stage writes gpurun_out/<tag>_<stage>.* ; a failing stage does not stop the later ones.
# Stages: tests | tests_fast | smoke | bench | bench_short | bench_ref | fri | fri_g9 | fri_diag | kat | mf | e2e<K> | launches | launches_tc | ncu_merkle | ncu_coop | ncu_ntt |
#         ncu_tc | ncu_tc2 | tc | tc_ab | tc_bench | tc_group | tc_batch | tc_streams (TCLIBS = tuning builds made with SMW_BUILD_SO / SMW_EXTRA_FLAGS) | sanitizer | umma | mgpu<N> | scale<N>
set -u
TAG=${1:?tag}; shift
OUT=gpurun_out; mkdir -p $OUT
export PYTHONUNBUFFERED=1
NCU_LIST="ncu --metrics gpu__time_duration.sum --clock-control none --csv"
NCU_FULL="ncu --set full --clock-control none --import-source on"
for stage in "$@"; do
  echo "=== stage $stage ($(date +%T))"
  case $stage in
    tests)      timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 --tb=short --durations=15 > $OUT/${TAG}_pytest_gpu.log 2>&1; tail -40 $OUT/${TAG}_pytest_gpu.log ;;
    tests_fast) timeout 900 python -m pytest tests -m gpu -q --maxfail=25 --tb=short --durations=10 --deselect tests/test_gpu_full_size.py > $OUT/${TAG}_pytest_gpu_fast.log 2>&1; tail -40 $OUT/${TAG}_pytest_gpu_fast.log ;;
    smoke)      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${TAG}_smoke.log 2>&1; tail -3 $OUT/${TAG}_smoke.log ;;
    bench)      timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; tail -c 3000 $OUT/${TAG}_bench.json; tail -5 $OUT/${TAG}_bench.err ;;
    bench_ref)  timeout 600 python bench.py --impl reference > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; cat $OUT/${TAG}_bench_reference.json ;;
    fri)        SMW_TRACE=1 timeout 600 python tools/fri_bench.py > $OUT/${TAG}_fri.json 2> $OUT/${TAG}_fri.err; cat $OUT/${TAG}_fri.json; tail -5 $OUT/${TAG}_fri.err ;;
    kat)        timeout 900 python -m pytest tests/test_gpu_proof_bytes_kat.py -m gpu -q --durations=10 > $OUT/${TAG}_kat.log 2>&1; tail -15 $OUT/${TAG}_kat.log ;;
    mf)         timeout 600 python tools/mf_bench.py > $OUT/${TAG}_mf.json 2> $OUT/${TAG}_mf.err; cat $OUT/${TAG}_mf.json; tail -5 $OUT/${TAG}_mf.err ;;
    launches)   timeout 900 $NCU_LIST -c 600 --log-file $OUT/${TAG}_ncu_launches.csv python bench.py --steps 1 --warmup 1 --log-n 16 --cols 16 --no-cpu --no-e2e > $OUT/${TAG}_ncu_launches.out 2>&1; tail -3 $OUT/${TAG}_ncu_launches.out ;;
    launches_fri) timeout 900 $NCU_LIST -c 900 --log-file $OUT/${TAG}_ncu_launches_fri.csv python tools/fri_bench.py --once > $OUT/${TAG}_ncu_launches_fri.out 2>&1; tail -3 $OUT/${TAG}_ncu_launches_fri.out ;;
    ncu_merkle) timeout 900 $NCU_FULL -k regex:k_merkle_level52 -c 1 -o $OUT/${TAG}_ncu_merkle_level52 -f python bench.py --steps 1 --warmup 0 --log-n 16 --cols 16 --no-cpu --no-e2e --no-fri > $OUT/${TAG}_ncu_merkle.out 2>&1
                ncu -i $OUT/${TAG}_ncu_merkle_level52.ncu-rep --page details --csv > $OUT/${TAG}_ncu_merkle_level52_details.csv 2>/dev/null; tail -3 $OUT/${TAG}_ncu_merkle.out ;;
    ncu_coop)   timeout 900 $NCU_FULL -k regex:k_nodes_coop52 -c 2 -o $OUT/${TAG}_ncu_nodes_coop52 -f python bench.py --steps 1 --warmup 0 --log-n 16 --cols 16 --no-cpu --no-e2e --no-fri > $OUT/${TAG}_ncu_coop.out 2>&1
                ncu -i $OUT/${TAG}_ncu_nodes_coop52.ncu-rep --page details --csv > $OUT/${TAG}_ncu_nodes_coop52_details.csv 2>/dev/null; tail -3 $OUT/${TAG}_ncu_coop.out ;;
    ncu_ntt)    timeout 900 $NCU_FULL -k regex:k_ntt -c 3 -o $OUT/${TAG}_ncu_ntt -f python bench.py --steps 1 --warmup 0 --log-n 18 --cols 8 --no-cpu --no-e2e --no-fri > $OUT/${TAG}_ncu_ntt.out 2>&1
                ncu -i $OUT/${TAG}_ncu_ntt.ncu-rep --page details --csv > $OUT/${TAG}_ncu_ntt_details.csv 2>/dev/null; tail -3 $OUT/${TAG}_ncu_ntt.out ;;
    sanitizer)  for tool in memcheck racecheck; do
                  timeout 1200 compute-sanitizer --tool $tool --error-exitcode 1 python -m pytest tests/test_gpu_sanitizer_subset.py -m gpu -q -x > $OUT/${TAG}_sanitizer_$tool.log 2>&1; echo "$tool exit $?" >> $OUT/${TAG}_sanitizer_$tool.log; tail -6 $OUT/${TAG}_sanitizer_$tool.log
                done ;;
    mgpu*)      N=${stage#mgpu}; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py > $OUT/${TAG}_multi_gpu_check_n$N.json 2> $OUT/${TAG}_multi_gpu_check_n$N.err; cat $OUT/${TAG}_multi_gpu_check_n$N.json; tail -8 $OUT/${TAG}_multi_gpu_check_n$N.err ;;
    scale*)     N=${stage#scale}; timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N > $OUT/${TAG}_bench_n$N.json 2> $OUT/${TAG}_bench_n$N.err; tail -c 2500 $OUT/${TAG}_bench_n$N.json; tail -5 $OUT/${TAG}_bench_n$N.err ;;

    ntt_ab)     timeout 600 python -m pytest tests/test_gpu_ntt_lde.py tests/test_gpu_full_size.py tests/test_gpu_ntt_distributed.py -m gpu -q --tb=short > $OUT/${TAG}_ntt_tests.log 2>&1; tail -5 $OUT/${TAG}_ntt_tests.log
                for v in 29 32; do SMW_NTT=$v timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_NTT=$v', 'value', d['value'], 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_ntt_ab.txt; done ;;
    ntt_thr)    for v in 128 64 32; do SMW_NTT_THREADS=$v timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_NTT_THREADS=$v', 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_ntt_threads.txt; done ;;
    tc)         timeout 600 python -m pytest tests/test_gpu_poseidon_tc.py -m gpu -q -x --tb=short > $OUT/${TAG}_tc_tests.log 2>&1; tail -25 $OUT/${TAG}_tc_tests.log ;;
    tc_ab)      for v in 0 16384; do SMW_TC_MIN=$v timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_TC_MIN=$v', 'value', d['value'], 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_tc_ab.txt; done ;;
    launches_tc) SMW_TC_MIN=4096 timeout 900 $NCU_LIST -c 400 --log-file $OUT/${TAG}_ncu_launches_tc.csv python bench.py --steps 1 --warmup 0 --log-n 16 --cols 16 --no-cpu --no-e2e --no-fri > $OUT/${TAG}_ncu_launches_tc.out 2>&1; tail -3 $OUT/${TAG}_ncu_launches_tc.out
                python - <<PYEOF | tee $OUT/${TAG}_ncu_launches_tc_summary.txt
import csv, collections, re
rows = [r for r in csv.reader(open("$OUT/${TAG}_ncu_launches_tc.csv")) if len(r) > 10]
h = rows[0]; ki = h.index("Kernel Name"); vi = h.index("Metric Value"); ui = h.index("Metric Unit")
d = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try: v = float(r[vi].replace(",", ""))
    except ValueError: continue
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    k = re.sub(r"\(.*", "", r[ki]); d[k][0] += 1; d[k][1] += v
tot = sum(v[1] for v in d.values())
for k, v in sorted(d.items(), key=lambda kv: -kv[1][1]): print("%-60s n=%4d  %9.3f ms  %5.1f%%" % (k[:60], v[0], v[1], 100 * v[1] / tot))
PYEOF
                ;;
    ncu_tc)     SMW_TC_MIN=4096 timeout 900 $NCU_FULL -k regex:k_tc_stage -s 1 -c 2 -o $OUT/${TAG}_ncu_tc_stage -f python bench.py --steps 1 --warmup 0 --log-n 16 --cols 16 --no-cpu --no-e2e --no-fri > $OUT/${TAG}_ncu_tc.out 2>&1
                ncu -i $OUT/${TAG}_ncu_tc_stage.ncu-rep --page details --csv > $OUT/${TAG}_ncu_tc_stage_details.csv 2>/dev/null; tail -3 $OUT/${TAG}_ncu_tc.out ;;
    tc_bench)   SMW_TC_MIN=${TCMIN:-16384} timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('TC', 'value', d['value'], 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_tc_bench.txt ;;
    tc_group)   for v in 8 4 2; do SMW_TC_GROUP=$v timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e --no-interp 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_TC_GROUP=$v', 'value', d['value'], 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_tc_group.txt; done ;;
    tc_batch)   for v in ${TCBATCHES:-1048576 524288 262144 131072}; do SMW_TC_BATCH=$v timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e --no-interp 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_TC_BATCH=$v', 'value', d['value'], 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_tc_batch.txt; done ;;
    fri_g9)     for v in 6 5 4; do SMW_TC_GROUP9=$v timeout 300 python tools/fri_bench.py 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_TC_GROUP9=$v', {k: (round(x['ms'], 2), x['class_ms']) for k, x in d.items()})" | tee -a $OUT/${TAG}_fri_g9.txt; done
                SMW_TC_MIN=0 timeout 300 python tools/fri_bench.py 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('SMW_TC_MIN=0 (interpreter)', {k: (round(x['ms'], 2), x['class_ms']) for k, x in d.items()})" | tee -a $OUT/${TAG}_fri_g9.txt ;;
    bench_short) timeout 600 python bench.py --no-fri --no-cpu --no-interp > $OUT/${TAG}_bench_short.json 2> $OUT/${TAG}_bench_short.err; python -c "import sys,json; d=json.loads(open('$OUT/${TAG}_bench_short.json').read().strip().splitlines()[-1]); print('value', d['value'], 'ms', d['ms_per_step'], 'e2e', d['e2e'])" ;;
    fri_diag)   SMW_TRACE=1 timeout 300 python tools/fri_diag.py > $OUT/${TAG}_fri_diag.json 2> $OUT/${TAG}_fri_diag.err; cat $OUT/${TAG}_fri_diag.json; grep -v "^\[smw\] \(build_forest\|fri_free\)" $OUT/${TAG}_fri_diag.err | tail -60 ;;
    tc_streams) for lib in ${TCLIBS:-libsmw.so}; do for v in ${TCSTREAMS:-1 2}; do SMW_LIB=$PWD/stark_mlwe_b200/$lib SMW_TC_STREAMS=$v timeout 300 python bench.py --steps 2 --warmup 1 --no-fri --no-cpu --no-e2e --no-interp 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib SMW_TC_STREAMS=$v', 'value', d['value'], 'ms', d['ms_per_step'], d['roofline']['step_share_ms'])" | tee -a $OUT/${TAG}_tc_streams.txt; done; done
                SMW_LIB=$PWD/stark_mlwe_b200/${TESTLIB:-libsmw.so} timeout 600 python -m pytest tests/test_gpu_poseidon_tc.py tests/test_gpu_full_size.py tests/test_gpu_fri_transcript.py -m gpu -q -x --tb=short > $OUT/${TAG}_tc_streams_tests.log 2>&1; tail -4 $OUT/${TAG}_tc_streams_tests.log ;;
    ncu_tc2)    for kn in k_tc_chain k_tc_stageILi0; do
                  SMW_TC_MIN=4096 SMW_TC_STREAMS=1 timeout 600 $NCU_FULL --kernel-name-base mangled -k regex:$kn -s 3 -c 1 -o $OUT/${TAG}_ncu_$kn -f python bench.py --steps 1 --warmup 0 --log-n 16 --cols 16 --no-cpu --no-e2e --no-fri --no-interp > $OUT/${TAG}_ncu_$kn.out 2>&1
                  ncu -i $OUT/${TAG}_ncu_$kn.ncu-rep --page details --csv > $OUT/${TAG}_ncu_${kn}_details.csv 2>/dev/null
                  ncu -i $OUT/${TAG}_ncu_$kn.ncu-rep --page raw --csv > $OUT/${TAG}_ncu_${kn}_raw.csv 2>/dev/null; tail -2 $OUT/${TAG}_ncu_$kn.out | cut -c1-200
                done ;;
    umma)       timeout 120 tools/_bin/umma_i8 ${UMMA_TILES:-16} $OUT/${TAG}_umma_dump.bin > $OUT/${TAG}_umma_i8.json 2> $OUT/${TAG}_umma_i8.err; echo "exit $?"; cat $OUT/${TAG}_umma_i8.json; tail -3 $OUT/${TAG}_umma_i8.err
                timeout 120 python tools/umma_i8_check.py $OUT/${TAG}_umma_dump.bin | tee $OUT/${TAG}_umma_check.json; rm -f $OUT/${TAG}_umma_dump.bin ;;
    diag26)     timeout 1200 python tools/diag_2p26.py 24 25 26 > $OUT/${TAG}_diag_2p26.json 2> $OUT/${TAG}_diag_2p26.err; cat $OUT/${TAG}_diag_2p26.json; tail -5 $OUT/${TAG}_diag_2p26.err ;;
    e2e*)       K=${stage#e2e}; timeout 1700 python tools/deep_fri_e2e.py $K > $OUT/${TAG}_deep_fri_e2e_2p$K.json 2> $OUT/${TAG}_deep_fri_e2e_2p$K.err; cat $OUT/${TAG}_deep_fri_e2e_2p$K.json; tail -5 $OUT/${TAG}_deep_fri_e2e_2p$K.err ;;
    *) echo "unknown stage $stage" ;;
  esac
done
echo "=== done ($(date +%T))"
