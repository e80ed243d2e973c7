#!/bin/bash
# Full GPU visit: pipe microbenchmark, all GPU tests, bench (default run + whole-batch verification), reference arm,
# secondary configs.  usage: tools/gpu_full.sh TAG
TAG=${1:-full}
O=gpurun_out
mkdir -p $O
./tools/ubench_pipes $O/${TAG}_ubench_pipes.json > $O/${TAG}_ubench_pipes.txt 2>&1; cat $O/${TAG}_ubench_pipes.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -3 $O/${TAG}_bench.err
timeout 900 python bench.py --steps 3 --warmup 3 --no-e2e --verify all > $O/${TAG}_bench_verify_all.json 2>> $O/${TAG}_bench.err; echo "verify-all rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2>> $O/${TAG}_bench.err; echo "ref rc=$?"
timeout 1500 python bench_configs.py erasure bch bch31 load --out $O/${TAG}_configs.jsonl > /dev/null 2> $O/${TAG}_configs.err; echo "configs rc=$?"; tail -3 $O/${TAG}_configs.err
ls -la $O | tail -12
