#!/bin/bash
# round 2, GPU call G: full -m gpu suite (new: allocation-free entry point, interval-filtered slab test), the no-alloc test alone in a
# fresh process, kernel probe with and without the slab filter (A/B library built with -DRTB_SLAB_FILTER=0)
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2g}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x -rfE --durations=5 > gpurun_out/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -15 gpurun_out/${tag}_gpu_tests_all.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu -rfE > gpurun_out/${tag}_gpu_tests_no_alloc_alone.log 2>&1; echo "no-alloc alone rc=$?"; tail -5 gpurun_out/${tag}_gpu_tests_no_alloc_alone.log
timeout 200 python tools/kernel_probe.py 128 0 > gpurun_out/${tag}_kernel_probe_filter.log 2>&1; tail -9 gpurun_out/${tag}_kernel_probe_filter.log
if [ -f ray-tracing_b200/librt_b200_nofilter.so ]; then
  RTB_CORE_LIB=librt_b200_nofilter.so timeout 200 python tools/kernel_probe.py 128 0 > gpurun_out/${tag}_kernel_probe_nofilter.log 2>&1; tail -9 gpurun_out/${tag}_kernel_probe_nofilter.log
fi
