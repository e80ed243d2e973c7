#!/bin/bash
# multi-GPU visit (gpurun --gpus N): the threads split on hardware, raw copy rates, e2e at 1..N devices.
# usage: tools/gpu_multi.sh TAG NCW_PER_GPU
TAG=${1:-multi}; NCW=${2:-4194304}
O=gpurun_out
mkdir -p $O
nvidia-smi -L | tee $O/${TAG}_gpus.txt
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1
lscpu | grep -E "NUMA|Socket|Model name|^CPU\(s\)" | tee $O/${TAG}_cpu.txt
free -g | head -2 | tee -a $O/${TAG}_cpu.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
timeout 600 python tools/pcie_probe.py --gb 4 > $O/${TAG}_pcie.jsonl 2> $O/${TAG}_pcie.err; echo "pcie rc=$?"; cat $O/${TAG}_pcie.jsonl
NG=$(nvidia-smi -L | wc -l)
for g in 1 2 4 8; do
  [ $g -le $NG ] || break
  timeout 900 python bench.py --split threads --gpus $g --ncw $NCW --e2e-steps 2 >> $O/${TAG}_threads.jsonl 2>> $O/${TAG}_threads.err; echo "threads $g rc=$?"
done
for g in 1 2 4 8; do
  [ $g -le $NG ] || break
  timeout 900 python bench_configs.py load5 --gpus $g --reps 2 --out $O/${TAG}_config5.jsonl > /dev/null 2>> $O/${TAG}_config5.err; echo "config5 $g rc=$?"
done
python - <<PY
import json
for l in open("$O/${TAG}_threads.jsonl"):
    d = json.loads(l); print(d["n_gpus"], round(d["value"],1), "GB/s", round(d["ms_per_step"],1), "ms", d["phases"])
for l in open("$O/${TAG}_config5.jsonl"):
    d = json.loads(l); print(d["gpus"], d["workload"][-6:], round(d["decode_ms"],2), "ms", round(d["payload_GBps"],1), "GB/s cpu", round(d["cpu"]["payload_GBps"],3))
PY
