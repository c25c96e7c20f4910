#!/bin/bash
# round 2, GPU call E: the whole -m gpu suite on the final code, the default bench line, then SMALL ncu captures (every ncu run under
# its own short timeout, reduced spp for the kernels whose full-size launch takes > 1 s: a wedged box is a strike)
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2e}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1200 python -m pytest tests -q -m gpu -rfE --durations=10 > gpurun_out/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -22 gpurun_out/${tag}_gpu_tests_all.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/${tag}_bench_default.json 2> gpurun_out/${tag}_bench_default.err; echo "bench rc=$?"; tail -c 600 gpurun_out/${tag}_bench_default.json; echo; tail -3 gpurun_out/${tag}_bench_default.err
timeout 200 python tools/kernel_probe.py 128 1 > gpurun_out/${tag}_kernel_probe.log 2>&1; tail -14 gpurun_out/${tag}_kernel_probe.log
cap() {  # name scene spp synth size kernel-regex skip [env]
  local name=$1 s=$2 n=$3 sn=$4 sz=$5 k=$6 skip=$7
  timeout 240 ncu --set full --clock-control none --import-source on --kernel-name regex:$k --launch-skip $skip --launch-count 1 -f -o gpurun_out/${tag}_$name \
      python bench.py --scene $s --spp $n --synth-n ${sn:-1000} ${sz:+--size $sz} --steps 1 --warmup 1 --profile > gpurun_out/${tag}_ncu_full_$name.log 2>&1
  echo "cap $name rc=$?"; ls -la gpurun_out/${tag}_$name.ncu-rep 2>&1 | tail -1
}
cap next_week_final_spp16 next_week_final 16 "" "" renderKernel 3
RTB_WAVEFRONT=0 cap cornell_renderKernel_spp32 cornell 32 "" "" renderKernel 3
cap cornell_wfTraverse_spp64 cornell 64 "" "" wfTraverseKernel 150
cap cornell_wfShade_spp64 cornell 64 "" "" wfShadeKernel 150
cap synthetic300_coop synthetic 1 300 480x270 renderKernelCoop 2
cap synthetic1000_coop synthetic 1 1000 240x135 renderKernelCoop 2
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${tag}_launches_cornell_spp64.csv \
    python bench.py --scene cornell --spp 64 --steps 1 --warmup 1 --profile > gpurun_out/${tag}_ncu_launches_cornell.log 2>&1; echo "cornell launch list rc=$?"
# summaries and the traffic table are made here, the reports themselves (8-14 MB each) stay on the box: gpurun_out/ is capped at 64 MiB
for f in gpurun_out/${tag}_*.ncu-rep; do python tools/ncu_summary.py $f 64 > ${f%.ncu-rep}.txt 2>&1; done
python tools/make_traffic.py gpurun_out/${tag}_traffic.json next_week_final_spp16=400,400,16,gpurun_out/${tag}_next_week_final_spp16.ncu-rep \
  cornell_renderKernel_spp32=500,500,32,gpurun_out/${tag}_cornell_renderKernel_spp32.ncu-rep cornell_wfTraverse_round=500,500,64,gpurun_out/${tag}_cornell_wfTraverse_spp64.ncu-rep \
  cornell_wfShade_round=500,500,64,gpurun_out/${tag}_cornell_wfShade_spp64.ncu-rep synthetic_300=480,270,1,gpurun_out/${tag}_synthetic300_coop.ncu-rep \
  synthetic_1000=240,135,1,gpurun_out/${tag}_synthetic1000_coop.ncu-rep > gpurun_out/${tag}_traffic.log 2>&1
rm -f gpurun_out/${tag}_*.ncu-rep
ls -la gpurun_out/ | grep ${tag}
