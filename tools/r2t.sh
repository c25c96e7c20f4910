#!/bin/bash
set -u
export PYTHONUNBUFFERED=1
mkdir -p gpurun_out
timeout 300 python tools/prec_speed_probe.py > gpurun_out/r2t_prec_speed.log 2>&1; cat gpurun_out/r2t_prec_speed.log
