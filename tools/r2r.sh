#!/bin/bash
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2r}
mkdir -p gpurun_out
timeout 400 python tools/knob_other_scenes.py 64 > gpurun_out/${tag}_knob_other_scenes.log 2>&1; cat gpurun_out/${tag}_knob_other_scenes.log | cut -c1-400
timeout 200 python tools/kernel_probe.py 128 0 > gpurun_out/${tag}_kernel_probe.log 2>&1; tail -9 gpurun_out/${tag}_kernel_probe.log | cut -c1-200
