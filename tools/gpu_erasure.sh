#!/bin/bash
# erasure-mode visit: parity tests that involve erasures, then configs[2] (16 Mi) with and without the erasure-only kernel.
TAG=${1:-era}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q -k "erasure or golden or quirk or config2" > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/${TAG}_pytest.log
timeout 900 python bench_configs.py erasure --out $O/${TAG}_erasure.jsonl > /dev/null 2> $O/${TAG}_erasure.err; echo "erasure rc=$?"; tail -3 $O/${TAG}_erasure.err
PP_RS255_NO_EE_KERNEL=1 timeout 900 python bench_configs.py erasure --out $O/${TAG}_erasure_old.jsonl > /dev/null 2>> $O/${TAG}_erasure.err; echo "erasure (old path) rc=$?"
python - <<PY
import json
for f in ("$O/${TAG}_erasure.jsonl", "$O/${TAG}_erasure_old.jsonl"):
    for l in open(f):
        d = json.loads(l); print(f, round(d["ms"], 2), "ms", d.get("kernels"), d.get("verify"))
PY
