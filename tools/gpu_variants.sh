#!/bin/bash
# A/B on the GPU box: per-kernel decode times (tools/dec_probe.py, 16 Mi codewords) of every library variant under
# build/variants/ against the in-tree library.  usage: tools/gpu_variants.sh TAG [ncw]
TAG=${1:-var}; NCW=${2:-16777216}
O=gpurun_out; mkdir -p $O
cp libpoporon_b200/lib/libpoporon_cuda.so /tmp/base_cuda.so
echo "== base" | tee $O/${TAG}_variants.txt
PYTHONPATH=. timeout 300 python tools/dec_probe.py $NCW 2>&1 | tail -4 | tee -a $O/${TAG}_variants.txt
for d in build/variants/*/; do
  n=$(basename $d)
  cp $d/libpoporon_cuda.so libpoporon_b200/lib/libpoporon_cuda.so
  echo "== $n" | tee -a $O/${TAG}_variants.txt
  PYTHONPATH=. timeout 300 python tools/dec_probe.py $NCW 2>&1 | tail -4 | tee -a $O/${TAG}_variants.txt
done
cp /tmp/base_cuda.so libpoporon_b200/lib/libpoporon_cuda.so
