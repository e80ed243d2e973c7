#!/bin/bash
# one ncu --set full capture with a launch skip.  usage: tools/gpu_ncu2.sh TAG KERNEL_REGEX SKIP COUNT [bench args...]
TAG=$1; RE=$2; SKIP=$3; CNT=$4; shift 4
O=gpurun_out
mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$RE" --launch-skip $SKIP -c $CNT \
  -o $O/${TAG} -f python bench.py --ncw 4194304 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline "$@" > $O/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
tail -2 $O/${TAG}_ncu.log | cut -c1-300
