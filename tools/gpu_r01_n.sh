#!/bin/bash
# long runs of the multiplier micro-benchmarks with nvidia-smi sampling: is the clock drop under a saturated IMAD.WIDE stream a power cap?
mkdir -p gpurun_out
nvidia-smi --query-gpu=timestamp,clocks.sm,clocks.max.sm,power.draw,power.limit,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.active,temperature.gpu --format=csv,noheader -lms 100 > gpurun_out/n_smi.csv &
SMI=$!
sleep 1
./tools/dfma_peak 600000 > gpurun_out/n_dfma_peak_long.jsonl 2>&1
sleep 1
kill $SMI
cat gpurun_out/n_dfma_peak_long.jsonl
awk -F, '{print $2,$4,$6,$9}' gpurun_out/n_smi.csv | sort | uniq -c | sort -rn | head -30
