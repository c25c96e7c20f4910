mkdir -p gpurun_out
B="python bench.py --spp 64 --steps 2 --warmup 3 --no-cpu-baseline"
for s in random random_book1 cornell next_week_final; do $B --scene $s > gpurun_out/v11_$s.json 2>&1; done
for rm in 3 10; do for s in cornell next_week_final; do RTB_RECT_MIN=$rm $B --scene $s > gpurun_out/v11_${s}_rm$rm.json 2>&1; done; done
