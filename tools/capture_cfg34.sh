set -u
for s in cornell next_week_final; do
  n=32; [ $s = next_week_final ] && n=16
  timeout 200 ncu --set full --clock-control none --import-source on --kernel-name regex:renderKernel --launch-skip 3 --launch-count 1 -f -o gpurun_out/r1g_$s python bench.py --scene $s --spp $n --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/r1g_ncu_$s.log 2>&1
  echo "$s rc=$?"; ls -la gpurun_out/r1g_$s.ncu-rep
done
