#!/bin/bash
# One GPU visit: parity tests, bench, ncu launch list, one ncu --set full capture (4 Mi codewords per launch).
# usage: tools/gpu_r02.sh TAG   (outputs under gpurun_out/TAG_*)
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -3 $O/${TAG}_pytest.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -c 3000 $O/${TAG}_bench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'rs255_fft_kernel|rs255_solve_kernel|rs255_fft_compact' -c 6 \
  -o $O/${TAG}_full -f python bench.py --ncw 4194304 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $O
