#!/bin/bash
# One GPU call at the end of a session: the newest tests first, then the default bench line, then the whole -m gpu suite.
# Every step has its own timeout and writes under gpurun_out/ as it goes, so a call that is cut short still leaves evidence.
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r1f}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 400 python -m pytest tests/test_gpu_quirks.py tests/test_gpu_progressive.py -q -m gpu -rfE --durations=10 > gpurun_out/${tag}_new_tests.log 2>&1
echo "new tests rc=$?"; tail -25 gpurun_out/${tag}_new_tests.log
timeout 300 python bench.py --steps 2 --warmup 3 > gpurun_out/${tag}_bench_random.json 2> gpurun_out/${tag}_bench_random.err
echo "bench rc=$?"; tail -c 700 gpurun_out/${tag}_bench_random.json; echo
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -rfE --durations=15 > gpurun_out/${tag}_parity_tests.log 2>&1
echo "parity tests rc=$?"; tail -30 gpurun_out/${tag}_parity_tests.log
