#!/bin/bash
# GPU call: full suite + kernel probe of the current build (tag = name of the variant under test)
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2i}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x -rfE > gpurun_out/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -6 gpurun_out/${tag}_gpu_tests_all.log
timeout 200 python tools/kernel_probe.py 128 ${2:-0} > gpurun_out/${tag}_kernel_probe.log 2>&1; tail -14 gpurun_out/${tag}_kernel_probe.log
