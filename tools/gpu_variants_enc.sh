#!/bin/bash
# A/B on the GPU box for encoder changes: parity tests of the encoder + tools/enc_probe.py for the in-tree library and
# every variant under build/variants/.  usage: tools/gpu_variants_enc.sh TAG [ncw]
TAG=${1:-venc}; NCW=${2:-16777216}
O=gpurun_out; mkdir -p $O
cp libpoporon_b200/lib/libpoporon_cuda.so /tmp/base_cuda.so
echo "== base" | tee $O/${TAG}_variants.txt
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q -k "golden or layout or encode or no_alloc or static or 239" 2>&1 | tail -2 | tee -a $O/${TAG}_variants.txt
PYTHONPATH=. timeout 300 python tools/enc_probe.py $NCW 2>&1 | tail -4 | tee -a $O/${TAG}_variants.txt
for d in build/variants/*/; do
  n=$(basename $d)
  cp $d/libpoporon_cuda.so libpoporon_b200/lib/libpoporon_cuda.so
  echo "== $n" | tee -a $O/${TAG}_variants.txt
  PYTHONPATH=. timeout 300 python tools/enc_probe.py $NCW 2>&1 | tail -4 | tee -a $O/${TAG}_variants.txt
done
cp /tmp/base_cuda.so libpoporon_b200/lib/libpoporon_cuda.so
