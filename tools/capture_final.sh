#!/bin/bash
# Round-end evidence for profiles/: per scene, (1) the bench line, (2) the ncu launch list of the same command, (3) one
# `ncu --set full` capture of renderKernel at the bench workload.  Each ncu pass runs only after the plain command exited 0.
# usage: tools/capture_final.sh TAG "scene[:spp] ..."
set -u
tag=$1; shift
for sc in $1; do
  s=${sc%%:*}; n=""; [ "$sc" != "$s" ] && n="--spp ${sc##*:}"
  python bench.py --scene $s $n --steps 2 --warmup 3 > gpurun_out/${tag}_bench_$s.json 2> gpurun_out/${tag}_bench_$s.err || { echo "bench $s failed"; tail -5 gpurun_out/${tag}_bench_$s.err; continue; }
  tail -c 600 gpurun_out/${tag}_bench_$s.json; echo
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_$s.csv \
      python bench.py --scene $s $n --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_launches_$s.log 2>&1
  ncu --set full --clock-control none --import-source on --kernel-name regex:renderKernel --launch-skip 3 --launch-count 1 -f -o gpurun_out/${tag}_$s \
      python bench.py --scene $s $n --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu_full_$s.log 2>&1
  ls -la gpurun_out/${tag}_$s.ncu-rep
done
