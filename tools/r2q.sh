#!/bin/bash
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2q}
mkdir -p gpurun_out
for v in "" _pf1 _pf2 _pf4; do
  RTB_CORE_LIB=librt_b200$v.so timeout 200 python tools/coop_probe.py >> gpurun_out/${tag}_coop_prefetch_ab.log 2>&1
done
cat gpurun_out/${tag}_coop_prefetch_ab.log
RTB_CORE_LIB=librt_b200_pf1.so timeout 300 python -m pytest tests/test_gpu_coop.py -q -m gpu -x > gpurun_out/${tag}_coop_tests_pf1.log 2>&1; echo "coop tests rc=$?"; tail -3 gpurun_out/${tag}_coop_tests_pf1.log
