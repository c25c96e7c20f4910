#!/bin/bash
# round 2, GPU call H: suite on the build with hoisted plane quotients, kernel probe, one source-level capture of cfg2 at 32 spp to count
# how often the slab filter hands over to the refined test
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2h}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -q -m gpu -x -rfE > gpurun_out/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -6 gpurun_out/${tag}_gpu_tests_all.log
timeout 200 python tools/kernel_probe.py 128 0 > gpurun_out/${tag}_kernel_probe.log 2>&1; tail -9 gpurun_out/${tag}_kernel_probe.log
for s in random next_week_final; do
timeout 240 ncu --set full --clock-control none --import-source on --kernel-name regex:renderKernel --launch-skip 3 --launch-count 1 -f -o gpurun_out/${tag}_$s \
  python bench.py --scene $s --spp 16 --steps 1 --warmup 1 --profile > gpurun_out/${tag}_ncu_$s.log 2>&1; echo "cap rc=$?"
python tools/ncu_summary.py gpurun_out/${tag}_$s.ncu-rep 64 > gpurun_out/${tag}_$s.txt 2>&1
done
rm -f gpurun_out/${tag}_*.ncu-rep
