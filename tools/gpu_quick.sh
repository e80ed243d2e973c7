#!/bin/bash
# quick GPU visit: parity tests, smoke, short bench.  usage: tools/gpu_quick.sh TAG
TAG=${1:-q}
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/${TAG}_pytest.log
timeout 600 python __graft_entry__.py smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -5 $O/${TAG}_smoke.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -c 2500 $O/${TAG}_bench.json; tail -5 $O/${TAG}_bench.err
