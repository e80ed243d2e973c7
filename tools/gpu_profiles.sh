#!/bin/bash
# Profile set of a round: bench (contract run), reference arm, ncu launch list of a short bench, one ncu --set full
# capture of every hot kernel (4 Mi codewords per launch).  usage: tools/gpu_profiles.sh TAG
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${TAG}_bench_reference.json 2>> $O/${TAG}_bench.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-e2e --no-cpu-baseline --verify none > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
# encode (fft, parity, rows) + the first real decode (screen, compact, solve, correct): skip the clean-decode passes
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'rs255_fft_kernel|rs255_parity_kernel|rs255_rows_kernel' -c 4 -o $O/${TAG}_full_a -f \
  python bench.py --ncw 4194304 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --verify none > $O/${TAG}_ncu_full_a.log 2>&1; echo "ncu full a rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on \
  -k regex:'rs255_solve_kernel|rs255_correct_kernel' --launch-skip 12 -c 2 -o $O/${TAG}_full_b -f \
  python bench.py --ncw 4194304 --steps 1 --warmup 0 --no-e2e --no-cpu-baseline --verify none > $O/${TAG}_ncu_full_b.log 2>&1; echo "ncu full b rc=$?"
ls -la $O | grep ${TAG}_
