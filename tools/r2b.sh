#!/bin/bash
# round 2, GPU call B: the new kernels and parity instruments: tests first (each file with its own timeout), then probes
set -u
export PYTHONUNBUFFERED=1
mkdir -p gpurun_out
for t in wavefront coop strict; do
  timeout 900 python -m pytest tests/test_gpu_$t.py -q -m gpu -rfE --durations=5 > gpurun_out/r2b_${t}_tests.log 2>&1; echo "$t tests rc=$?"; tail -25 gpurun_out/r2b_${t}_tests.log
done
timeout 600 python tools/kernel_probe.py 128 1 > gpurun_out/r2b_kernel_probe.log 2>&1; tail -16 gpurun_out/r2b_kernel_probe.log
timeout 300 python tools/parity_classify.py gpurun_out/r2b_parity_classification_default.json > gpurun_out/r2b_classify_default.log 2>&1; tail -30 gpurun_out/r2b_classify_default.log
timeout 300 python tools/parity_classify.py gpurun_out/r2b_parity_classification_strict.json --strict > gpurun_out/r2b_classify_strict.log 2>&1; tail -12 gpurun_out/r2b_classify_strict.log
timeout 300 python tools/parity_fractions.py gpurun_out/r2b_parity_fractions_strict.json --strict > gpurun_out/r2b_fractions_strict.log 2>&1; tail -3 gpurun_out/r2b_fractions_strict.log
timeout 300 python tools/parity_fractions.py gpurun_out/r2b_parity_fractions_default.json > gpurun_out/r2b_fractions_default.log 2>&1; tail -3 gpurun_out/r2b_fractions_default.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/r2b_smoke.log
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -rfE -k "90k or converged or fp64_matches" --durations=8 > gpurun_out/r2b_parity_new.log 2>&1; echo "parity-new rc=$?"; tail -15 gpurun_out/r2b_parity_new.log
