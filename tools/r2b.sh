#!/bin/bash
# round 2, GPU call B: strict-math parity (new tests), classifier, strict parity fractions, smoke, 90k-sphere trace
set -u
export PYTHONUNBUFFERED=1
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_strict.py -q -m gpu -x -rfE --durations=8 > gpurun_out/r2b_strict_tests.log 2>&1; echo "strict tests rc=$?"; tail -30 gpurun_out/r2b_strict_tests.log
timeout 300 python tools/parity_classify.py gpurun_out/r2b_parity_classification_default.json > gpurun_out/r2b_classify_default.log 2>&1; tail -30 gpurun_out/r2b_classify_default.log
timeout 300 python tools/parity_classify.py gpurun_out/r2b_parity_classification_strict.json --strict > gpurun_out/r2b_classify_strict.log 2>&1; tail -12 gpurun_out/r2b_classify_strict.log
timeout 300 python tools/parity_fractions.py gpurun_out/r2b_parity_fractions_strict.json --strict > gpurun_out/r2b_fractions_strict.log 2>&1; tail -3 gpurun_out/r2b_fractions_strict.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/r2b_smoke.log
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -rfE -k "90k or converged or fp64_matches" --durations=8 > gpurun_out/r2b_parity_new.log 2>&1; echo "parity-new rc=$?"; tail -15 gpurun_out/r2b_parity_new.log
