#!/usr/bin/env python
"""Raw host<->device copy rates of the box: concurrent pinned H2D + D2H with plain cudaMemcpyAsync (torch copies on
two streams per device, one host thread per device), at 1/2/4/8 GPUs.  The bound bench.py's e2e is compared with.
usage: python tools/pcie_probe.py [--gb 2] [--chunk-mb 64]   -> one JSON line per device count"""
import argparse
import json
import threading
import time

import torch


def run(ndev, gb, chunk_mb, mode):
    chunk = chunk_mb << 20
    nchunks = max(1, int(gb * (1 << 30)) // chunk)
    bufs = []
    for d in range(ndev):
        torch.cuda.set_device(d)
        h_in = torch.empty(chunk * 2, dtype=torch.uint8).pin_memory()
        h_out = torch.empty(chunk * 2, dtype=torch.uint8).pin_memory()
        dv = torch.empty(chunk * 2, dtype=torch.uint8, device=f"cuda:{d}")
        bufs.append((h_in, h_out, dv, torch.cuda.Stream(device=d), torch.cuda.Stream(device=d)))
    barrier = threading.Barrier(ndev + 1)
    times = [0.0] * ndev

    def worker(d):
        torch.cuda.set_device(d)
        h_in, h_out, dv, s1, s2 = bufs[d]
        barrier.wait()
        t0 = time.perf_counter()
        for i in range(nchunks):
            o = (i & 1) * chunk
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    dv[o:o + chunk].copy_(h_in[o:o + chunk], non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    h_out[o:o + chunk].copy_(dv[o:o + chunk], non_blocking=True)
        s1.synchronize()
        s2.synchronize()
        times[d] = time.perf_counter() - t0

    th = [threading.Thread(target=worker, args=(d,)) for d in range(ndev)]
    for t in th:
        t.start()
    barrier.wait()
    for t in th:
        t.join()
    tmax = max(times)
    per_dir = ndev * nchunks * chunk / tmax / 1e9
    return per_dir


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=4.0, help="bytes per device and direction")
    ap.add_argument("--chunk-mb", type=int, default=64)
    args = ap.parse_args()
    total = torch.cuda.device_count()
    for ndev in (1, 2, 4, 8):
        if ndev > total:
            break
        run(ndev, 0.5, args.chunk_mb, "both")  # warm-up (pins, contexts)
        line = {"gpus": ndev, "chunk_mb": args.chunk_mb, "gb_per_device_per_direction": args.gb}
        for mode in ("h2d", "d2h", "both"):
            line[mode + "_GBps_per_direction"] = round(run(ndev, args.gb, args.chunk_mb, mode), 1)
        line["both_total_GBps"] = round(2 * line["both_GBps_per_direction"], 1)
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
