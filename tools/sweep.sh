mkdir -p gpurun_out
B="python bench.py --spp 64 --steps 2 --warmup 3 --no-cpu-baseline"
for sm in 16 24 33; do for s in random cornell next_week_final; do RTB_SHADE_MIN=$sm $B --scene $s > gpurun_out/v10_${s}_sm$sm.json 2>&1; done; done
timeout 300 python bench.py --scene synthetic --synth-n 1000 --steps 1 --warmup 3 > gpurun_out/v10_synthetic.json 2> gpurun_out/v10_synth.err
