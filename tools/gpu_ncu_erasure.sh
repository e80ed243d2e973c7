#!/bin/bash
# one ncu --set full capture (with source counters) of the erasure-only kernel on 4 Mi codewords.  usage: tools/gpu_ncu_erasure.sh TAG
TAG=${1:-era}
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rs255_erasure_kernel' --launch-skip 1 -c 1 \
  -o $O/${TAG}_erasure -f python bench_configs.py erasure --ncw 4194304 --reps 2 > $O/${TAG}_ncu_erasure.log 2>&1; echo "ncu rc=$?"; tail -2 $O/${TAG}_ncu_erasure.log
