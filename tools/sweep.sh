mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
B="python bench.py --spp 64 --steps 2 --warmup 3 --no-cpu-baseline"
for s in random random_book1 cornell next_week_final; do $B --scene $s > gpurun_out/v6_$s.json 2>&1; done
$B --scene random --fast > gpurun_out/v6_random_fast.json 2>&1
RTB_KEEP_NODE=12 RTB_KEEP_LEAF=8 $B --scene random > gpurun_out/v6_random_kn12kl8.json 2>&1
