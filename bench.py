#!/usr/bin/env python
"""bench.py - RS(255,223) batched encode + decode throughput (BASELINE.json metric) on N B200s.

Contract (see the task brief): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0.

Workload (BASELINE.json configs[1]): per GPU, 16 Mi RS(255,223) codewords as packed 255-byte records;
one step = encode all messages (223 B -> 32 B parity) + decode all corrupted codewords (0-16 random byte
errors each, uniform; positions over data and parity; corrected in place).  Inputs are resident in HBM
when the timed region starts (`value`); `e2e` is the same step through the host-pointer C API
(poporon_encode_batch / poporon_decode_batch on page-locked host buffers, copies inside the timed region).

Each GPU owns a contiguous slice of the codeword index space (weak scaling, no data-path collective:
codewords are independent - SURVEY 8e); torch.distributed is used only for the barrier and the
max-over-ranks of the elapsed time.

`--impl reference` times the reference's own CPU implementation (oracle/_ref/libpoporon_ref.so = the
unmodified reference compiled by oracle/Makefile; falls back to the oracle port when it is absent) with one
independent poporon_t per host core on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "rs255_223_encode_decode_payload_GBps"
UNIT = "GB/s"
K_MSG, K_PAR, K_CW = 223, 32, 255
SEED_MSG, SEED_ERR = 0x5EED0001, 0x5EED0002
L2_BYTES = 126 * 1024 * 1024


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--ncw", type=int, default=16 * 1024 * 1024, help="codewords per GPU per step")
    ap.add_argument("--emax", type=int, default=16, help="errors per codeword are uniform in [0, emax]")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample", type=int, default=0, help="codewords in the CPU baseline sample (0 = auto)")
    ap.add_argument("--mem-budget-gb", type=float, default=140.0)
    ap.add_argument("--split", choices=["ranks", "threads"], default="ranks",
                    help="ranks: one process per GPU (torchrun, the contract line).  threads: ONE process, one host "
                         "thread per device inside libpoporon (poporon_use_devices), host-resident batch, no "
                         "torch.distributed / NCCL - the arrangement north_star describes")
    ap.add_argument("--verify", choices=["sample", "all", "none"], default="sample",
                    help="untimed comparison with the CPU reference after the timed loop: a 1 Mi prefix + a 64 Ki "
                         "strided sample of the batch (default), the whole batch, or nothing")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def alu_peak(sm_mhz):
    """Integer-ALU roof in lane-operations per second: the LOP3 rate measured by tools/ubench_pipes.cu
    (profiles/r02_ubench_pipes.json, warp-instructions per clock and SM at the attribute clock) x 32 lanes x SMs x
    the SM clock sampled during the timed region."""
    path = os.path.join(ROOT, "profiles", "r02_ubench_pipes.json")
    lanes, sms, src = 64.0, 148, "fallback: 64 lanes/clk/SM (DESIGN.md section 4)"
    if os.path.exists(path):
        with open(path) as f:
            u = json.load(f)
        lanes, sms, src = float(u["pipes"]["LOP3 xor3"]["lanes_per_clk_per_sm"]), int(u["sms"]), "profiles/r02_ubench_pipes.json (LOP3 xor3)"
    return lanes * sms * sm_mhz * 1e6, src


# Useful integer-ALU lane-operations per codeword of each ALU-bound kernel: the operation count of its formulation
# with nothing but the GF(2) arithmetic counted (DESIGN.md section 4 derives every term).
ALU_FLOOR = {
    # BM: 304 coefficient steps (sum over the 32 iterations r of ceil(r / 2) + 1: the longest register among a warp's
    # 1024 codewords sets the trip count) x (2 x 64 on the held multiples of gamma and Delta + 8 select + 64 for the
    # next discrepancy) + 32 x (2 x 21 multiples + 28 reduction); evaluator 9.2 K; Chien search as a transform 17 x
    # (2 x 244 DFT + 192 multipliers + 16 + 120 zero tests) + 136 basis change; Forney on 16 list slots x (23 Horner steps
    # x 64 on the root's held multiples + 2 x 21 for the multiples + 18 squaring + 2 products x 92 + a^254 = 4 x 92 + 7 x
    # 18); 28 bit-matrix transposes x 256 - per 32 queued codewords
    "rs255_solve_kernel": (304 * 200 + 32 * 70 + 9152 + 17 * 816 + 136 + 16 * 2210 + 7168) / 32.0,
    # transform 18598 + 64 word-columns x (256 transpose + 32 realign) + 255 x 8 basis change, per 32 codewords
    "rs255_fft_kernel<SCREEN>": (18598 + 64 * 288 + 255 * 8) / 32.0,
    "rs255_fft_kernel<ENCODE>": (18598 + 56 * 288 + 223 * 8) / 32.0,
    # 32 syndromes x (two subset tables 22 + 256 plane updates), per 32 codewords
    "rs255_parity_kernel": 32 * (22 + 256) / 32.0,
    "rs255_rows_kernel": 8 * 256 / 32.0,
}


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per codeword for each kernel, from the committed
    `ncu --set full` capture summary (profiles/ncu_traffic.json; written by tools/ncu_traffic.py)."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return {}
    with open(path) as f:
        return json.load(f)


# ----------------------------------------------------------------------------------------- multi-rank plumbing
def shard_of(rank, ncw_per_rank):
    """The contiguous slice of the global codeword index space a rank works on (weak scaling: every rank holds
    ncw_per_rank codewords; the workload generators are pure functions of the global index, so any rank can produce
    any slice).  Returns (first global index, count)."""
    return rank * ncw_per_rank, ncw_per_rank


def max_over_ranks(values, device=None):
    """Element-wise maximum of a list of floats over all ranks (identity without torch.distributed): the job's time is
    the slowest rank's.  Works on any backend (nccl under torchrun on GPUs, gloo in the CPU tests)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(v) for v in values]
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


# ----------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region (pynvml)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thr = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksThrottleReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        if self.nv:
            self._thr = threading.Thread(target=self._loop, daemon=True)
            self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join()
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ----------------------------------------------------------------------------------------- CPU arms
def cpu_workload(n, emax, seed=1234):
    """Synthetic codewords of the same distribution as the GPU workload, generated with numpy."""
    import numpy as np

    rng = np.random.default_rng(seed)
    msg = rng.integers(0, 256, (n, K_MSG), dtype=np.uint8)
    e = rng.integers(0, emax + 1, n)
    order = np.argsort(rng.random((n, K_CW)), axis=1)[:, :max(emax, 1)]
    mags = rng.integers(1, 256, (n, max(emax, 1)), dtype=np.uint8)
    return msg, e, order, mags


def time_cpu(lib, msg, e, order, mags, threads):
    """One CPU step on the sample: encode then decode; returns seconds."""
    import numpy as np

    n = len(msg)
    t0 = time.perf_counter()
    par = lib.rs_encode(msg, threads=threads)
    t1 = time.perf_counter()
    cw = np.concatenate([msg, par], 1)
    rows = np.arange(n)
    for k in range(order.shape[1]):
        sel = e > k
        cw[rows[sel], order[sel, k]] ^= mags[sel, k]
    d, p = np.ascontiguousarray(cw[:, :K_MSG]), np.ascontiguousarray(cw[:, K_MSG:])
    t2 = time.perf_counter()
    out = lib.rs_decode(d, p, threads=threads)
    t3 = time.perf_counter()
    assert out[2].all(), "CPU baseline failed to decode a correctable codeword"
    return (t1 - t0) + (t3 - t2)


def load_cpu_lib():
    from oraclelib import Oracle, have_reference

    if have_reference():
        return Oracle("reference"), "reference"
    return Oracle("port"), "port"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    lib, kind = load_cpu_lib()
    cores = os.cpu_count() or 1
    sample = args.cpu_sample or min(args.ncw, cores * 20000)
    msg, e, order, mags = cpu_workload(sample, args.emax)
    for _ in range(min(args.warmup, 1)):
        time_cpu(lib, msg[: max(1, sample // 8)], e[: max(1, sample // 8)], order[: max(1, sample // 8)],
                 mags[: max(1, sample // 8)], cores)
    total = 0.0
    for _ in range(args.steps):
        total += time_cpu(lib, msg, e, order, mags, cores)
    per_step = total / max(args.steps, 1)
    value = 2 * sample * K_MSG / per_step / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": workload_config(args, sample_per_step=sample),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"{sample} codewords per step (encode + decode), one handle per thread"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, **extra):
    cfg = {"workload": "RS(255,223) batched encode+decode, 0-%d random byte errors per codeword "
                       "(BASELINE configs[1])" % args.emax,
           "codewords_per_gpu": args.ncw, "record_bytes": K_CW, "layout": "packed 255-byte records",
           "errors": f"uniform 0..{args.emax} per codeword, distinct positions over data+parity",
           "cache": "inputs larger than L2 (%.2f GB per pass vs 126 MB L2)" % (args.ncw * K_CW / 1e9)}
    cfg.update(extra)
    return cfg


# ----------------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import libpoporon_b200.api as api

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # NCCL prints its version banner on stdout at the first collective: keep stdout for the ONE JSON line
        sys.stdout.flush()
        keep = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(keep, 1)
            os.close(keep)

    n = args.ncw
    i0, _ = shard_of(rank, n)  # this rank's slice of the global codeword index space
    K, W = args.steps, args.warmup
    rec_bytes = n * K_CW
    nbuf = max(1, min(K + W, int(args.mem_budget_gb * 1e9 // rec_bytes) - 1))
    reinject = nbuf < K + W
    stream = torch.cuda.current_stream().cuda_stream

    h = api.Poporon.rs_default()
    assert h.use_devices([local])

    clean = torch.empty(rec_bytes, dtype=torch.uint8, device=dev)
    api.synth_fill(local, clean.data_ptr(), K_CW, K_MSG, n, SEED_MSG, i0, stream)
    assert h.encode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, n, stream)
    bufs, dirty = [], []

    def corrupt(b):
        """XOR buffer b's deterministic error pattern in (a second application would remove it again)"""
        api.synth_inject(local, bufs[b].data_ptr(), K_CW, K_CW, n, SEED_ERR + b, i0, 0, 0, args.emax, 0, stream)
        dirty[b] = not dirty[b]

    for b in range(nbuf):
        bufs.append(clean.clone())
        dirty.append(False)
        corrupt(b)
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    cor = torch.empty(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()

    def step(s, ev=None):
        buf = bufs[s % nbuf]
        if not dirty[s % nbuf]:  # only when K + W exceeds the buffer ring (reported as reinject_in_timed_region)
            corrupt(s % nbuf)
        dirty[s % nbuf] = False  # the decode below restores it
        if ev:
            ev[0].record()
        assert h.encode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, n, stream)
        if ev:
            ev[1].record()
        assert h.decode_batch_async(buf.data_ptr(), K_CW, K_MSG, buf.data_ptr() + K_MSG, K_CW, n, ok.data_ptr(),
                                    cor.data_ptr(), stream=stream)
        if ev:
            ev[2].record()

    for s in range(W):
        step(s)
    torch.cuda.synchronize()
    # correctness gate on the last warm-up buffer: everything decodable, restored to the clean records
    if W > 0:
        last = bufs[(W - 1) % nbuf]
        assert bool(ok.all().item()), "a correctable codeword failed to decode"
        assert torch.equal(last, clean), "decoded records differ from the encoded records"

    # clean-decode pass (north_star: "encode and clean-decode against HBM"): decode of the uncorrupted records =
    # screen + compaction only; separate, untimed for `value`, averaged over 5 passes with CUDA events.
    # The 4.28 GB record buffer is >> L2 and the passes before it touch other buffers.
    cev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
    h.decode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, n, ok.data_ptr(),
                         cor.data_ptr(), stream=stream)
    for i in range(6):
        cev[i].record()
        if i < 5:
            h.decode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, n, ok.data_ptr(),
                                 cor.data_ptr(), stream=stream)
    torch.cuda.synchronize()
    clean_ms = cev[0].elapsed_time(cev[5]) / 5
    assert bool(ok.all().item()) and int(cor.sum().item()) == 0, "clean records did not decode as clean"

    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(K)]
    sampler = ClockSampler(local)
    launches0 = h.launches
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    t_start = torch.cuda.Event(enable_timing=True)
    t_stop = torch.cuda.Event(enable_timing=True)
    h.profile(True)  # CUDA event pairs around every kernel of the handle, on the launching stream
    t_start.record()
    for s in range(K):
        step(W + s, evs[s])
    t_stop.record()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    kprof = h.profile_read()
    h.profile(False)
    if world > 1:
        dist.barrier()
    launches = h.launches - launches0
    elapsed_ms = t_start.elapsed_time(t_stop)
    enc_ms = sum(e[0].elapsed_time(e[1]) for e in evs) / K
    dec_ms = sum(e[1].elapsed_time(e[2]) for e in evs) / K
    elapsed_ms, enc_ms, dec_ms = max_over_ranks([elapsed_ms, enc_ms, dec_ms], dev)
    ms_per_step = elapsed_ms / K
    payload_step = 2 * n * K_MSG * world
    value = payload_step / (ms_per_step * 1e-3) / 1e9

    # mean errors per codeword (algorithmic write bytes of the decode)
    mean_err = float(cor.float().mean().item())
    peak, peak_src = peaks()
    enc_alg = n * (K_MSG + K_PAR)
    dec_alg = n * (K_CW + 5 + mean_err)
    # roofline of the dominant kernel: algorithmic bytes per launch / its average launch duration (CUDA events
    # around each launch inside the timed region, see poporon_cuda_rs_profile)
    dirty_frac = float((cor > 0).float().mean().item())
    alg_bytes = {  # per launch, DESIGN.md section 4
        # the encoder is two kernels: message rows -> syndrome planes (223 B read), planes -> parity rows (32 B written);
        # the 32 B per codeword of planes between them are implementation traffic, not algorithmic bytes
        "rs255_fft_kernel<ENCODE>": n * K_MSG,
        "rs255_parity_kernel": n * K_PAR,
        "rs255_rows_kernel": n * K_PAR,
        # (with the zero-syndrome exit fused in: + status 5 B per codeword + a queue entry per dirty one)
        "rs255_fft_kernel<SCREEN>": n * (K_CW + 0.125 + 5 + (32 + 4) * dirty_frac),
        "rs255_compact_kernel": n * (0.25 + 5 + 4 * dirty_frac),  # (only when the status arrays are not 16-byte aligned)
        # queue entry + syndrome row + status per dirty codeword; error record (2 B) + read-modify-write of the
        # corrected byte (2 B) per correction (the solver applies its own corrections)
        "rs255_solve_kernel": n * dirty_frac * (4 + 32 + 5) + n * mean_err * (2 + 2),
        "rs255_correct_kernel": 8,
    }
    sm_mhz = float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0) if isinstance(clocks, dict) else 1965.0
    apeak, apeak_src = alu_peak(sm_mhz)
    units = {"rs255_solve_kernel": n * dirty_frac}  # codewords a launch works on (default: the whole batch)
    kernels = {}
    for name, (cnt, ms) in kprof.items():
        avg = ms / cnt
        kernels[name] = {"launches_per_step": cnt / K, "avg_ms": avg, "share_of_step": ms / elapsed_ms,
                         "alg_GBps": alg_bytes.get(name, 0) / (avg * 1e-3) / 1e9}
        if name in ALU_FLOOR:
            ops = ALU_FLOOR[name] * units.get(name, n)
            kernels[name].update({"alu_floor_lane_ops_per_codeword": ALU_FLOOR[name],
                                  "alu_Tlaneops": ops / (avg * 1e-3) / 1e12,
                                  "alu_frac": ops / (avg * 1e-3) / apeak})
    traffic = ncu_traffic()
    dom = max(kernels, key=lambda k: kernels[k]["avg_ms"] * kernels[k]["launches_per_step"]) if kernels else None
    if dom:
        tr = traffic.get(dom)
        hbm = {"achieved_GBps": kernels[dom]["alg_GBps"], "bytes_per_launch": alg_bytes.get(dom, 0), "peak_GBps": peak,
               "frac": kernels[dom]["alg_GBps"] / peak, "peak_source": peak_src,
               "traffic": tr["bytes_per_codeword"] * n if tr else None}
        if dom in ALU_FLOOR:
            # error-heavy decode is bound by the integer ALU pipe (north_star): useful lane-operations per second
            # against the measured LOP3 rate; the HBM view of the same kernel stays beside it
            roof = {"kernel": dom, "bound": "alu", "achieved": kernels[dom]["alu_Tlaneops"], "peak": apeak / 1e12,
                    "unit": "Tlane-op/s", "frac": kernels[dom]["alu_frac"], "peak_source": apeak_src,
                    "floor_lane_ops_per_codeword": ALU_FLOOR[dom], "codewords_per_launch": units.get(dom, n),
                    "avg_launch_ms": kernels[dom]["avg_ms"], "share_of_step": kernels[dom]["share_of_step"],
                    "traffic": hbm["traffic"], "hbm": hbm}
        else:
            roof = {"kernel": dom, "bound": "hbm", "achieved": hbm["achieved_GBps"], "peak": peak, "unit": "GB/s",
                    "frac": hbm["frac"], "peak_source": peak_src, "bytes_per_launch": hbm["bytes_per_launch"],
                    "avg_launch_ms": kernels[dom]["avg_ms"], "share_of_step": kernels[dom]["share_of_step"],
                    "traffic": hbm["traffic"]}
        if tr:
            roof["traffic_source"] = tr["source"]
    else:
        roof = {"kernel": "rs decode (whole phase)", "bound": "hbm", "achieved": dec_alg / (dec_ms * 1e-3) / 1e9,
                "traffic": None, "peak": peak, "unit": "GB/s", "peak_source": peak_src}
        roof["frac"] = roof["achieved"] / peak
    phases = {"encode_ms": enc_ms, "decode_ms": dec_ms,
              "encode_payload_GBps": n * K_MSG / (enc_ms * 1e-3) / 1e9,
              "decode_payload_GBps": n * K_MSG / (dec_ms * 1e-3) / 1e9,
              "encode_hbm_frac": enc_alg / (enc_ms * 1e-3) / 1e9 / peak,
              "decode_hbm_frac": dec_alg / (dec_ms * 1e-3) / 1e9 / peak,
              "clean_decode_ms": clean_ms,
              "clean_decode_payload_GBps": n * K_MSG / (clean_ms * 1e-3) / 1e9,
              "clean_decode_hbm_frac": n * (K_CW + 5.25) / (clean_ms * 1e-3) / 1e9 / peak,
              "dirty_fraction": dirty_frac}

    # ---- e2e: host-pointer API on page-locked buffers, copies inside the timed region -------------
    e2e = None
    if not args.no_e2e:
        hmsg = torch.empty(n * K_MSG, dtype=torch.uint8).pin_memory()
        hpar = torch.empty(n * K_PAR, dtype=torch.uint8).pin_memory()
        hrec = torch.empty(rec_bytes, dtype=torch.uint8).pin_memory()
        hok = torch.empty(n, dtype=torch.uint8).pin_memory()
        hcor = torch.empty(n, dtype=torch.int32).pin_memory()
        hmsg.view(n, K_MSG).copy_(clean.view(n, K_CW)[:, :K_MSG])
        e2e_steps = max(1, args.e2e_steps)
        tot = 0.0
        for s in range(e2e_steps + 1):  # first pass is a warm-up (allocates the staging slots)
            src = bufs[s % nbuf]
            if not dirty[s % nbuf]:  # decoded in place during the timed loop: corrupt it again (untimed)
                corrupt(s % nbuf)
            hrec.copy_(src)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            assert h.encode_batch(hmsg.data_ptr(), K_MSG, K_MSG, hpar.data_ptr(), K_PAR, n)
            assert h.decode_batch(hrec.data_ptr(), K_CW, K_MSG, hrec.data_ptr() + K_MSG, K_CW, n, hok.data_ptr(),
                                  hcor.data_ptr())
            dt = time.perf_counter() - t0
            if s > 0:
                tot += dt
        assert bool(hok.all().item())
        assert torch.equal(hrec.view(n, K_CW)[:, K_MSG:], hpar.view(n, K_PAR)), "e2e parity mismatch"
        tmax = torch.tensor([tot / e2e_steps], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        e2e_s = float(tmax.item())
        e2e = {"value": payload_step / e2e_s / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": world * n * (K_MSG + K_CW), "d2h_bytes_per_step": world * n * (K_PAR + K_CW + 5),
               "ms_per_step": e2e_s * 1e3, "steps": e2e_steps,
               "api": "poporon_encode_batch + poporon_decode_batch on page-locked host buffers"}

    # ---- CPU baseline next to it (rank 0, N = 1 only) ------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        lib, kind = load_cpu_lib()
        cores = os.cpu_count() or 1
        sample = args.cpu_sample or min(n, cores * 20000)
        rec = clean.view(n, K_CW)[:sample].cpu().numpy()
        if not dirty[0]:
            corrupt(0)
        torch.cuda.synchronize()
        badh = bufs[0].view(n, K_CW)[:sample].cpu().numpy()
        t0 = time.perf_counter()
        par = lib.rs_encode(np.ascontiguousarray(rec[:, :K_MSG]), threads=cores)
        out = lib.rs_decode(np.ascontiguousarray(badh[:, :K_MSG]), np.ascontiguousarray(badh[:, K_MSG:]),
                            threads=cores)
        dt = time.perf_counter() - t0
        assert (par == rec[:, K_MSG:]).all() and out[2].all() and (out[0] == rec[:, :K_MSG]).all(), \
            "CPU baseline and GPU disagree on the sample"
        cpu = {"value": 2 * sample * K_MSG / dt / 1e9, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"first {sample} codewords of the GPU batch (encode + decode), one handle per thread; "
                         "outputs compared with the GPU's"}

    # ---- oracle comparison at size (untimed, rank 0, N = 1): data, parity, ok and corrected_num of a 1 Mi prefix
    # plus a 64 Ki strided sample of the 16 Mi batch - or of all of it - against the compiled reference ------------
    verify = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.verify != "none":
        lib, kind = load_cpu_lib()
        cores = os.cpu_count() or 1
        if args.verify == "all":
            idx = torch.arange(n, device=dev)
        else:
            pre = min(n, 1 << 20)
            stride = max(1, n // 65536)
            idx = torch.unique(torch.cat([torch.arange(pre, device=dev), torch.arange(0, n, stride, device=dev)]))
        if not dirty[0]:
            corrupt(0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        mism = 0
        # in slices of 1 Mi codewords: the received words of the slice, the GPU's decode of the WHOLE buffer, the
        # reference's decode of the slice
        rx = bufs[0].view(n, K_CW)
        parts = [idx[i:i + (1 << 20)] for i in range(0, idx.numel(), 1 << 20)]
        rxs = [rx[p_].cpu().numpy() for p_ in parts]
        assert h.decode_batch_async(bufs[0].data_ptr(), K_CW, K_MSG, bufs[0].data_ptr() + K_MSG, K_CW, n, ok.data_ptr(),
                                    cor.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        dirty[0] = False
        for p_, r_ in zip(parts, rxs):
            got = rx[p_].cpu().numpy()
            gok = ok[p_].cpu().numpy()
            gcor = cor[p_].cpu().numpy().astype(np.uint32)
            wd, wp, wok, wcor = lib.rs_decode(np.ascontiguousarray(r_[:, :K_MSG]), np.ascontiguousarray(r_[:, K_MSG:]),
                                              threads=cores)
            mism += int(((got[:, :K_MSG] != wd).any(1) | (got[:, K_MSG:] != wp).any(1) | (gok != wok) |
                         (gcor != wcor)).sum())
        verify = {"codewords": int(idx.numel()), "mismatches": mism, "against": kind, "mode": args.verify,
                  "fields": "data, parity, ok, corrected_num", "seconds": round(time.perf_counter() - t0, 1)}
        assert mism == 0, f"GPU decode differs from the CPU {kind} on {mism} codewords"

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": workload_config(args, corrupted_buffers=nbuf, reinject_in_timed_region=reinject,
                                      mean_errors_per_codeword=mean_err),
            "roofline": roof, "phases": phases, "kernels": kernels, "clocks": clocks, "gpu_launches": int(launches),
        }
        if e2e:
            line["e2e"] = e2e
        if cpu:
            line["cpu_baseline"] = cpu
        if verify:
            line["verify"] = verify
        print(json.dumps(line), flush=True)
    h.close()
    if world > 1:
        dist.destroy_process_group()


def run_threads(args):
    """north_star's multi-GPU arrangement: one process, the batch split contiguously over the devices by the library
    (one host thread per device, csrc/host/pipeline.c), no collective.  Host-resident data only, so `value` IS the
    end-to-end number (copies inside the timed region)."""
    import numpy as np
    import torch

    import libpoporon_b200.api as api

    N = args.gpus
    if torch.cuda.device_count() < N:
        raise SystemExit(f"--split threads --gpus {N}: only {torch.cuda.device_count()} devices visible")
    n_per, n = args.ncw, args.ncw * N
    h = api.Poporon.rs_default()
    assert h.use_devices(list(range(N)))
    hmsg = torch.empty(n * K_MSG, dtype=torch.uint8).pin_memory()
    hpar = torch.empty(n * K_PAR, dtype=torch.uint8).pin_memory()
    hrec = torch.empty(n * K_CW, dtype=torch.uint8).pin_memory()
    bad = torch.empty(n * K_CW, dtype=torch.uint8)  # the corrupted records, copied into hrec before every step
    hok = torch.empty(n, dtype=torch.uint8).pin_memory()
    hcor = torch.empty(n, dtype=torch.int32).pin_memory()
    # workload: generated on device 0 slice by slice (same generators and seeds as the ranks split)
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream().cuda_stream
    gen = api.Poporon.rs_default()
    want_par = torch.empty(n * K_PAR, dtype=torch.uint8)
    for r in range(N):
        rec = torch.empty(n_per * K_CW, dtype=torch.uint8, device=dev)
        api.synth_fill(0, rec.data_ptr(), K_CW, K_MSG, n_per, SEED_MSG, r * n_per, stream)
        assert gen.encode_batch_async(rec.data_ptr(), K_CW, K_MSG, rec.data_ptr() + K_MSG, K_CW, n_per, stream)
        torch.cuda.synchronize()
        v = rec.view(n_per, K_CW)
        hmsg.view(n, K_MSG)[r * n_per:(r + 1) * n_per].copy_(v[:, :K_MSG])
        want_par.view(n, K_PAR)[r * n_per:(r + 1) * n_per].copy_(v[:, K_MSG:])
        api.synth_inject(0, rec.data_ptr(), K_CW, K_CW, n_per, SEED_ERR, r * n_per, 0, 0, args.emax, 0, stream)
        torch.cuda.synchronize()
        bad.view(n, K_CW)[r * n_per:(r + 1) * n_per].copy_(v)
        del rec, v
    gen.close()
    torch.cuda.empty_cache()
    samplers = [ClockSampler(i) for i in range(N)]
    K, W = max(1, args.e2e_steps), 1
    tot = enc_tot = 0.0
    for s in range(W + K):
        hrec.copy_(bad)
        hpar.zero_()
        if s == W:
            for sp in samplers:
                sp.start()
            l0 = h.launches
        t0 = time.perf_counter()
        assert h.encode_batch(hmsg.data_ptr(), K_MSG, K_MSG, hpar.data_ptr(), K_PAR, n)
        t1 = time.perf_counter()
        assert h.decode_batch(hrec.data_ptr(), K_CW, K_MSG, hrec.data_ptr() + K_MSG, K_CW, n, hok.data_ptr(),
                              hcor.data_ptr())
        t2 = time.perf_counter()
        if s >= W:
            tot += t2 - t0
            enc_tot += t1 - t0
    clocks = [sp.stop() for sp in samplers]
    launches = h.launches - l0
    assert bool(hok.all().item()), "a correctable codeword failed to decode"
    assert torch.equal(hpar, want_par), "parity differs from the single-device encode"
    assert torch.equal(hrec.view(n, K_CW)[:, :K_MSG], hmsg.view(n, K_MSG)) and \
        torch.equal(hrec.view(n, K_CW)[:, K_MSG:], want_par.view(n, K_PAR)), "decoded records differ"
    per_step = tot / K
    value = 2 * n * K_MSG / per_step / 1e9
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": N, "steps": K, "warmup": W,
            "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic", "split": "threads (one process, one host thread per device, no NCCL)",
            "config": workload_config(args, corrupted_buffers=1, reinject_in_timed_region=False,
                                      mean_errors_per_codeword=float(hcor.float().mean().item())),
            "phases": {"encode_ms": enc_tot / K * 1e3, "decode_ms": (tot - enc_tot) / K * 1e3},
            "clocks": {"sm_mhz": min(c.get("sm_mhz") or 0 for c in clocks), "sm_max_mhz": clocks[0].get("sm_max_mhz"),
                       "reasons": sorted({r for c in clocks for r in c.get("reasons", [])})},
            "gpu_launches": int(launches),
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": n * (K_MSG + K_CW),
                    "d2h_bytes_per_step": n * (K_PAR + K_CW + 5), "ms_per_step": per_step * 1e3, "steps": K,
                    "api": "poporon_use_devices + poporon_encode_batch + poporon_decode_batch on page-locked host "
                           "buffers, ONE call over all devices"}}
    print(json.dumps(line), flush=True)
    h.close()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    elif args.split == "threads":
        run_threads(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
