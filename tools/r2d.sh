#!/bin/bash
# round 2, multi-GPU call: tools/r2d.sh N — in-library multi-GPU tests, then bench.py at N GPUs the way the driver launches it
set -u
export PYTHONUNBUFFERED=1
N=${1:-2}; tag=r2d_n$N
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -rfEs --durations=5 > gpurun_out/${tag}_multi_tests.log 2>&1; echo "multi tests rc=$?"; tail -15 gpurun_out/${tag}_multi_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 \
    > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/${tag}_bench.json; echo; tail -8 gpurun_out/${tag}_bench.err
