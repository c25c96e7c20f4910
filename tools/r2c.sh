#!/bin/bash
# round 2, GPU call C: the evidence for profiles/: the default bench line (all scenes), launch list of the same command, one
# `ncu --set full` capture of the dominant kernel per scene at the bench workload (after the plain command exited 0), traffic table
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2c}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 300 python tools/wf_sweep.py 128 > gpurun_out/${tag}_wf_sweep.log 2>&1; tail -32 gpurun_out/${tag}_wf_sweep.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/${tag}_bench_default.json 2> gpurun_out/${tag}_bench_default.err; echo "bench rc=$?"; tail -c 1500 gpurun_out/${tag}_bench_default.json; echo; tail -5 gpurun_out/${tag}_bench_default.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${tag}_bench_reference.json 2>&1; tail -c 400 gpurun_out/${tag}_bench_reference.json; echo
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches_default.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-scene > gpurun_out/${tag}_ncu_launches.log 2>&1; echo "launch list rc=$?"
cap() {  # scene spp synth size kernel-regex skip
  local s=$1 n=$2 sn=$3 sz=$4 k=$5 skip=$6
  timeout 900 ncu --set full --clock-control none --import-source on --kernel-name regex:$k --launch-skip $skip --launch-count 1 -f -o gpurun_out/${tag}_${s}${sn}_$k \
      python bench.py --scene $s --spp $n --synth-n ${sn:-1000} ${sz:+--size $sz} --steps 1 --warmup 1 --no-cpu-baseline --no-per-scene > gpurun_out/${tag}_ncu_full_${s}${sn}_$k.log 2>&1
  ls -la gpurun_out/${tag}_${s}${sn}_$k.ncu-rep
}
cap random 1024 "" "" renderKernel 2
cap random_book1 1024 "" "" renderKernel 2
cap next_week_final 256 "" "" renderKernel 2
cap cornell 1000 "" "" wfTraverseKernel 300
cap cornell 1000 "" "" wfShadeKernel 300
RTB_WAVEFRONT=0 cap cornell 1000 "" "" renderKernel 2
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${tag}_launches_cornell.csv \
    python bench.py --scene cornell --spp 100 --steps 1 --warmup 1 --no-cpu-baseline --no-per-scene > gpurun_out/${tag}_ncu_launches_cornell.log 2>&1
cap synthetic 1 300 480x270 renderKernelCoop 2
cap synthetic 1 1000 240x135 renderKernelCoop 2
for f in gpurun_out/${tag}_*.ncu-rep; do python tools/ncu_summary.py $f 64 > ${f%.ncu-rep}.txt 2>&1; done
ls -la gpurun_out/
timeout 1500 python -m pytest tests -q -m gpu -rfE --durations=10 -x > gpurun_out/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -25 gpurun_out/${tag}_gpu_tests_all.log
