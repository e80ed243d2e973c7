#!/bin/bash
# A/B on the GPU box for erasure-mode changes: configs[2] (bench_configs.py erasure, verified against the reference) for the
# in-tree library and every variant under build/variants/.  usage: tools/gpu_variants_era.sh TAG
TAG=${1:-vera}
O=gpurun_out; mkdir -p $O
cp libpoporon_b200/lib/libpoporon_cuda.so /tmp/base_cuda.so
run() {
  timeout 600 python bench_configs.py erasure --out $O/${TAG}_$1.jsonl > /dev/null 2>> $O/${TAG}.err
  python - <<PY
import json
for l in open("$O/${TAG}_$1.jsonl"):
    d = json.loads(l); print("$1", round(d["ms"], 2), "ms", {k: round(v["avg_ms"], 2) for k, v in d["kernels"].items()}, "mismatches", d["verify"]["mismatches"])
PY
}
run base
for d in build/variants/*/; do
  n=$(basename $d)
  cp $d/libpoporon_cuda.so libpoporon_b200/lib/libpoporon_cuda.so
  run $n
done
cp /tmp/base_cuda.so libpoporon_b200/lib/libpoporon_cuda.so
