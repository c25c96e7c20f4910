#!/bin/bash
# round 2, GPU call A: measured issue peaks, observed parity fractions of the round-1 kernel, F32 vs MIXED throughput
set -u
export PYTHONUNBUFFERED=1
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
./tools/fp32_peak > gpurun_out/r2a_fp32_peak.json; cat gpurun_out/r2a_fp32_peak.json
timeout 300 python tools/prec_probe.py 128 > gpurun_out/r2a_prec_probe.log 2>&1; tail -12 gpurun_out/r2a_prec_probe.log
timeout 600 python tools/parity_fractions.py gpurun_out/r2a_parity_fractions_default.json > gpurun_out/r2a_parity_fractions.log 2>&1; tail -5 gpurun_out/r2a_parity_fractions.log
