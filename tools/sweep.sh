mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
B="python bench.py --spp 64 --steps 2 --warmup 3 --no-cpu-baseline"
for s in random random_book1 cornell next_week_final; do $B --scene $s > gpurun_out/v4_$s.json 2>&1; done
for kn in 8 12 20; do RTB_KEEP_NODE=$kn $B --scene random > gpurun_out/v4_random_kn$kn.json 2>&1; done
for kl in 6 9 16; do RTB_KEEP_LEAF=$kl $B --scene random > gpurun_out/v4_random_kl$kl.json 2>&1; done
$B --scene random --fast > gpurun_out/v4_random_fast.json 2>&1
