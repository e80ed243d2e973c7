#!/bin/bash
# Builds libpoporon_cuda.so variants that differ in the solver's compile-time switches, for A/B runs on the GPU box
# (tools/gpu_variants.sh).  usage: tools/solve_variants.sh NAME "-DFLAG=..." [NAME2 "-D..."] ...
set -e
cd "$(dirname "$0")/../libpoporon_b200/csrc/cuda"
ROOT=$(cd ../../.. && pwd)
make -s >/dev/null 2>&1
while [ $# -ge 2 ]; do
  NAME=$1; FLAGS=$2; shift 2
  OUT=$ROOT/build/variants/$NAME
  mkdir -p $OUT
  /usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a \
    -Xcompiler -fPIC -I$ROOT/include -I. $FLAGS -Xptxas -v -c ${FILE:-rs255}.cu -o $OUT/${FILE:-rs255}.o 2> $OUT/ptxas.txt
  grep -A2 "${KERNEL:-rs255_solve_kernel}" $OUT/ptxas.txt | grep -E "registers|spill" | sed "s/^/[$NAME] /"
  OBJS=$(ls $ROOT/build/cuda/*.o | grep -v -E "/(${FILE:-rs255}|synth)\.o")
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $OUT/libpoporon_cuda.so $OBJS $OUT/${FILE:-rs255}.o -cudart static
done
