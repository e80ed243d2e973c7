#!/bin/bash
# one ncu --set full capture (with source counters) of the solver on 4 Mi codewords.  usage: tools/gpu_ncu_solve.sh TAG
TAG=${1:-solve}
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rs255_solve_kernel' --launch-skip 6 -c 1 \
  -o $O/${TAG}_solve -f python bench.py --ncw 4194304 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --verify none \
  > $O/${TAG}_ncu_solve.log 2>&1; echo "ncu rc=$?"; tail -3 $O/${TAG}_ncu_solve.log
