#!/bin/bash
# one ncu --set full capture.  usage: tools/gpu_ncu.sh TAG KERNEL_REGEX [COUNT] [bench args...]
TAG=$1; RE=$2; CNT=${3:-4}; shift 3
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" -c $CNT \
  -o $O/${TAG} -f python bench.py --ncw 4194304 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline "$@" > $O/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 $O/${TAG}_ncu.log
