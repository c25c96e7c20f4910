#!/bin/bash
# short call: wavefront sort probe, cfg1 / cfg2 --set full captures at 128 spp (a 1024-spp launch x 40 replays does not fit a 300 s timeout),
# traffic metrics of ONE full-size launch for cfg1 / cfg2 / cfg3
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2u}
g=gpurun_out
mkdir -p $g
timeout 200 python tools/wf_sort_probe.py > $g/${tag}_wf_sort_probe.log 2>&1; cat $g/${tag}_wf_sort_probe.log
for s in random random_book1; do
  timeout 200 ncu --set full --clock-control none --import-source on --kernel-name regex:renderKernel --launch-skip 3 --launch-count 1 -f -o $g/${tag}_${s}_spp128 \
      python bench.py --scene $s --spp 128 --steps 1 --warmup 1 --profile > $g/${tag}_ncu_full_${s}_spp128.log 2>&1; echo "cap $s rc=$?"
  python tools/ncu_summary.py $g/${tag}_${s}_spp128.ncu-rep 64 > $g/${tag}_${s}_spp128.txt 2>&1
done
M=dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors.sum,l1tex__t_sectors.sum,gpu__time_duration.sum
for s in random random_book1 next_week_final; do
  timeout 200 ncu --metrics $M --clock-control none --kernel-name regex:renderKernel --launch-skip 3 --launch-count 1 -f -o $g/${tag}_traffic_$s \
      python bench.py --scene $s --steps 1 --warmup 1 --profile > $g/${tag}_ncu_traffic_$s.log 2>&1; echo "traffic $s rc=$?"
done
python tools/make_traffic.py $g/${tag}_traffic.json random=600,400,1024,$g/${tag}_traffic_random.ncu-rep random_book1=600,400,1024,$g/${tag}_traffic_random_book1.ncu-rep \
  next_week_final=400,400,256,$g/${tag}_traffic_next_week_final.ncu-rep random_spp128=600,400,128,$g/${tag}_random_spp128.ncu-rep \
  random_book1_spp128=600,400,128,$g/${tag}_random_book1_spp128.ncu-rep > $g/${tag}_traffic.log 2>&1; tail -3 $g/${tag}_traffic.log
rm -f $g/${tag}_*.ncu-rep
ls $g | grep ${tag}
