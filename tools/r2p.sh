#!/bin/bash
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2p}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_wavefront.py tests/test_gpu_multi.py tests/test_gpu_progressive.py -q -m gpu -x -rfE > gpurun_out/${tag}_gpu_tests_wf.log 2>&1; echo "wf tests rc=$?"; tail -4 gpurun_out/${tag}_gpu_tests_wf.log
timeout 400 python tools/wf_guide_sweep.py > gpurun_out/${tag}_wf_guide_sweep.log 2>&1; cat gpurun_out/${tag}_wf_guide_sweep.log
