#!/bin/bash
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2j}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x -rfE > gpurun_out/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -4 gpurun_out/${tag}_gpu_tests_all.log
RTB_CORE_LIB=librt_b200_dbg.so timeout 200 python tools/unsure_probe.py > gpurun_out/${tag}_unsure_probe.log 2>&1; cat gpurun_out/${tag}_unsure_probe.log
timeout 200 python tools/kernel_probe.py 128 0 > gpurun_out/${tag}_kernel_probe.log 2>&1; tail -9 gpurun_out/${tag}_kernel_probe.log
