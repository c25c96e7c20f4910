#!/bin/bash
# round 2, final single-GPU evidence on the final code: whole -m gpu suite, smoke, default bench line, reference arm, launch lists,
# ncu --set full captures (every ncu run under its own timeout; reduced spp where a full-size launch x 40 replays would take minutes),
# metrics-only traffic captures at FULL size for cfg3 / cfg4, summaries and the traffic table made on the box (reports stay there: 64 MiB cap)
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2z}
g=gpurun_out
mkdir -p $g
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1200 python -m pytest tests -q -m gpu -rfE --durations=8 > $g/${tag}_gpu_tests_all.log 2>&1; echo "all gpu tests rc=$?"; tail -14 $g/${tag}_gpu_tests_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $g/${tag}_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 $g/${tag}_smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > $g/${tag}_bench_default.json 2> $g/${tag}_bench_default.err; echo "bench rc=$?"; head -c 400 $g/${tag}_bench_default.json; echo; tail -3 $g/${tag}_bench_default.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $g/${tag}_bench_reference.json 2> $g/${tag}_bench_reference.err; echo "reference arm rc=$?"; head -c 300 $g/${tag}_bench_reference.json; echo
timeout 200 python tools/kernel_probe.py 128 1 > $g/${tag}_kernel_probe.log 2>&1; tail -14 $g/${tag}_kernel_probe.log | cut -c1-160
# launch lists (gpu__time_duration per launch, the recipe's pass): cfg2 full size, cfg4 at 64 spp
timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $g/${tag}_launches_random_spp1024.csv \
    python bench.py --scene random --steps 2 --warmup 1 --profile > $g/${tag}_ncu_launches_random.log 2>&1; echo "random launch list rc=$?"
timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $g/${tag}_launches_cornell_spp64.csv \
    python bench.py --scene cornell --spp 64 --steps 1 --warmup 1 --profile > $g/${tag}_ncu_launches_cornell.log 2>&1; echo "cornell launch list rc=$?"
cap() {  # name scene spp synth size kernel-regex skip
  local name=$1 s=$2 n=$3 sn=$4 sz=$5 k=$6 skip=$7
  timeout 300 ncu --set full --clock-control none --import-source on --kernel-name regex:$k --launch-skip $skip --launch-count 1 -f -o $g/${tag}_$name \
      python bench.py --scene $s --spp $n --synth-n ${sn:-1000} ${sz:+--size $sz} --steps 1 --warmup 1 --profile > $g/${tag}_ncu_full_$name.log 2>&1
  echo "cap $name rc=$?"; ls -la $g/${tag}_$name.ncu-rep 2>&1 | tail -1
}
cap random_spp1024 random 1024 "" "" renderKernel 3
cap random_book1_spp1024 random_book1 1024 "" "" renderKernel 3
cap next_week_final_spp16 next_week_final 16 "" "" renderKernel 3
cap cornell_wfTraverse_spp1000 cornell 1000 "" "" wfTraverseKernel 600
cap cornell_wfShade_spp1000 cornell 1000 "" "" wfShadeKernel 600
cap synthetic300_coop synthetic 1 300 480x270 renderKernelCoop 2
cap synthetic1000_coop synthetic 1 1000 240x135 renderKernelCoop 2
# traffic at FULL size, metrics only (4 metrics = 1-2 replays per launch): cfg3 one launch, cfg4 every launch of one frame
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,l1tex__t_sectors.sum,gpu__time_duration.sum
timeout 300 ncu --metrics $M --clock-control none --kernel-name regex:renderKernel --launch-skip 3 --launch-count 1 -f -o $g/${tag}_traffic_next_week_final \
    python bench.py --scene next_week_final --steps 1 --warmup 1 --profile > $g/${tag}_ncu_traffic_next_week_final.log 2>&1; echo "traffic cfg3 rc=$?"
timeout 400 ncu --metrics $M --clock-control none --kernel-name regex:wf -f -o $g/${tag}_traffic_cornell \
    python bench.py --scene cornell --steps 1 --warmup 0 --profile > $g/${tag}_ncu_traffic_cornell.log 2>&1; echo "traffic cfg4 rc=$?"
for f in $g/${tag}_*.ncu-rep; do case $f in *traffic*) ;; *) python tools/ncu_summary.py $f 64 > ${f%.ncu-rep}.txt 2>&1;; esac; done
python tools/make_traffic.py $g/${tag}_traffic.json random=600,400,1024,$g/${tag}_random_spp1024.ncu-rep random_book1=600,400,1024,$g/${tag}_random_book1_spp1024.ncu-rep \
  next_week_final=400,400,256,$g/${tag}_traffic_next_week_final.ncu-rep cornell=500,500,1000,$g/${tag}_traffic_cornell.ncu-rep,,sum \
  next_week_final_spp16=400,400,16,$g/${tag}_next_week_final_spp16.ncu-rep cornell_wfTraverse_round=500,500,1000,$g/${tag}_cornell_wfTraverse_spp1000.ncu-rep \
  cornell_wfShade_round=500,500,1000,$g/${tag}_cornell_wfShade_spp1000.ncu-rep synthetic_300=480,270,1,$g/${tag}_synthetic300_coop.ncu-rep \
  synthetic_1000=240,135,1,$g/${tag}_synthetic1000_coop.ncu-rep > $g/${tag}_traffic.log 2>&1
rm -f $g/${tag}_*.ncu-rep
ls -la $g/ | grep ${tag} | awk '{print $5, $9}'
