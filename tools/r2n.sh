#!/bin/bash
# A/B of the plane-quotient variants: default (cuboids + rects hoisted), rects literal, both literal; kernel probe x2 each
set -u
export PYTHONUNBUFFERED=1
tag=${1:-r2n}
mkdir -p gpurun_out
for v in "" _norect _noplane; do
  for rep in 1 2; do
    RTB_CORE_LIB=librt_b200$v.so timeout 200 python tools/kernel_probe.py 128 0 > gpurun_out/${tag}_kernel_probe${v:-_default}_$rep.log 2>&1
    echo "variant ${v:-default} rep $rep"; head -8 gpurun_out/${tag}_kernel_probe${v:-_default}_$rep.log | grep -v "random_book1.*wavefront\|random .*wavefront" 
  done
done
