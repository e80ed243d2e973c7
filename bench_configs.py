#!/usr/bin/env python
"""bench_configs.py - the secondary BASELINE.json workloads (configs[2], [3], [4]) on one B200.

`bench.py` carries the contract line (configs[1]).  This script times the other full-size configurations through
the same C ABI on device-resident data and prints one JSON line per workload (appended to --out when given):

  erasure   configs[2]: RS(255,223) erasure decoding, 16 Mi codewords, 0-32 sorted erasures + 0-3 unknown errors
            (the reference's slot-order quirk makes rows with unknown errors "succeed" with
            wrong bytes - the size-independent check here is on the rows WITHOUT unknown errors, which must be
            restored exactly; tests/test_gpu_parity.py compares every class with the oracle at small sizes)
  bch       configs[3]: BCH(15,5) t=3, 64 Mi codewords, 0-4 flipped bits each; b <= 3 must be corrected
  load      configs[4] (single-GPU leg): RS(255,223) decode of 128 Mi codewords at i.i.d. symbol-error rates
            0 %, 1 % and 5 % (uncorrectable codewords are part of the 5 % load); check: every codeword reported
            ok is a codeword again (second, clean decode of the corrected buffer: 0 corrections)

Besides the size-independent properties, a 256 Ki sample of configs[2] (rows WITH unknown errors and with
2v + c > 32 included) and of configs[4] at 5 % (14 % of it uncorrectable) is byte-compared - data, parity, ok,
corrected_num - with the CPU reference (oracle/_ref when present, else the pinned restatement), untimed.  Timing: CUDA events on the launching stream, one untimed warm-up on
a different buffer, inputs far larger than L2.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

K_MSG, K_PAR, K_CW = 223, 32, 255


def emit(args, line):
    s = json.dumps(line)
    print(s, flush=True)
    if args.out:
        with open(args.out, "a") as f:
            f.write(s + "\n")


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"])
    return 6650.0


def sample_index(torch, n, dev, total=256 * 1024):
    """a prefix of 3/4 of the sample plus a strided quarter over the whole batch"""
    pre = min(n, total * 3 // 4)
    stride = max(1, n // (total // 4))
    return torch.unique(torch.cat([torch.arange(pre, device=dev), torch.arange(0, n, stride, device=dev)]))


def oracle_compare(rx, got, ok, cor, mode=0, pos=None, cnt=None):
    """rx/got: [m, 255] numpy (received / decoded records).  Returns (mismatching rows, oracle ok fraction, kind)."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oraclelib import Oracle, have_reference

    kind = "reference" if have_reference() else "port"
    orc = Oracle(kind)
    wd, wp, wok, wcor = orc.rs_decode(np.ascontiguousarray(rx[:, :K_MSG]), np.ascontiguousarray(rx[:, K_MSG:]),
                                      mode=mode, pos=pos, cnt=cnt, threads=os.cpu_count() or 1)
    bad = (got[:, :K_MSG] != wd).any(1) | (got[:, K_MSG:] != wp).any(1) | (ok != wok) | (cor.astype(np.uint32) != wcor)
    return int(bad.sum()), float(wok.mean()), kind


def timed(torch, fn, reps):
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    for i in range(reps):
        ev[i].record()
        fn(i)
    ev[reps].record()
    torch.cuda.synchronize()
    return [ev[i].elapsed_time(ev[i + 1]) for i in range(reps)]


def run_erasure(args, torch, api, dev):
    n = args.ncw or 16 * 1024 * 1024
    stream = torch.cuda.current_stream().cuda_stream
    h = api.Poporon.rs_default()
    clean = torch.empty(n * K_CW, dtype=torch.uint8, device=dev)
    api.synth_fill(0, clean.data_ptr(), K_CW, K_MSG, n, 0x5EED0001, 0, stream)
    assert h.encode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, n, stream)
    reps = args.reps
    bufs, poss, cnts = [], [], []
    for b in range(reps + 1):
        buf = clean.clone()
        pos = torch.empty(n * 32, dtype=torch.int32, device=dev)
        cnt = torch.empty(n, dtype=torch.int32, device=dev)
        api.synth_erasures(0, buf.data_ptr(), K_CW, K_MSG, K_CW, n, 0x5EED0300 + b, pos.data_ptr(), 32,
                           cnt.data_ptr(), 0, 32, stream)
        bufs.append(buf), poss.append(pos), cnts.append(cnt)
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    cor = torch.empty(n, dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    dirty_rows = {reps - 1: (bufs[reps - 1].view(n, K_CW) != clean.view(n, K_CW)).any(1)}

    def dec(b):
        assert h.decode_batch_async(bufs[b].data_ptr(), K_CW, K_MSG, bufs[b].data_ptr() + K_MSG, K_CW, n,
                                    ok.data_ptr(), cor.data_ptr(), pos=poss[b].data_ptr(), pos_stride=32,
                                    cnt=cnts[b].data_ptr(), stream=stream)

    idx = sample_index(torch, n, dev)
    rx_s = bufs[reps - 1].view(n, K_CW)[idx].cpu().numpy()
    dec(reps)  # warm-up on the extra buffer
    torch.cuda.synchronize()
    h.profile(True)
    ms = timed(torch, dec, reps)
    prof = h.profile_read()
    h.profile(False)
    # property: rows the reference semantics restore exactly = rows whose only damage is on listed (sorted)
    # erasure positions.  Identify them without the oracle: after decode they equal the clean record.
    last = reps - 1
    restored = (bufs[last].view(n, K_CW) == clean.view(n, K_CW)).all(1)
    okb = ok.bool()
    frac_ok = float(okb.float().mean().item())
    frac_restored = float(restored.float().mean().item())
    # synth_erasures leaves 90 % of the rows without unknown errors: all of those must be restored and ok
    assert frac_restored > 0.895, frac_restored
    assert bool((okb | ~restored).all().item()), "a restored row was reported as failed"
    # oracle comparison of the sample, every outcome class included (unknown errors, 2v + c > 32, failures)
    import numpy as np

    pos_s = poss[last].view(n, 32)[idx].cpu().numpy().astype(np.uint32)
    cnt_s = cnts[last][idx].cpu().numpy().astype(np.uint32)
    mism, ork, kind = oracle_compare(rx_s, bufs[last].view(n, K_CW)[idx].cpu().numpy(), ok[idx].cpu().numpy(),
                                     cor[idx].cpu().numpy(), mode=1, pos=pos_s, cnt=cnt_s)
    assert mism == 0, f"erasure decode differs from the CPU {kind} on {mism} of {idx.numel()} sampled codewords"
    t = sum(ms) / len(ms)
    mean_cnt = float(cnts[last].float().mean().item())
    alg = n * (K_CW + 5 + 4 * mean_cnt + 4 + float(cor.float().mean().item()))
    emit(args, {"workload": "configs[2]: RS(255,223) erasure decode, 0-32 sorted erasures + 0-3 unknown errors",
                "codewords": n, "ms": t, "payload_GBps": n * K_MSG / t / 1e6, "alg_GBps": alg / t / 1e6,
                "hbm_frac": alg / t / 1e6 / hbm_peak(), "bound": "integer ALU / shared memory (general per-thread stage)",
                "dirty_fraction": float(dirty_rows[last].float().mean().item()), "ok_fraction": frac_ok,
                "restored_fraction": frac_restored, "mean_erasures": mean_cnt,
                "verify": {"codewords": int(idx.numel()), "mismatches": mism, "against": kind,
                           "oracle_ok_fraction": ork, "fields": "data, parity, ok, corrected_num"},
                "kernels": {k: {"launches": c, "avg_ms": m / c} for k, (c, m) in prof.items()}})
    h.close()


def run_bch(args, torch, api, dev):
    n = args.ncw or 64 * 1024 * 1024
    stream = torch.cuda.current_stream().cuda_stream
    h = api.Poporon.bch_default()
    g = torch.Generator(device=dev)
    g.manual_seed(0x5EED0400)
    data = torch.randint(0, 256, (n,), dtype=torch.uint8, device=dev, generator=g)  # top 3 bits = garbage
    reps = args.reps
    # distinct input and output buffers per pass (64 + 128 MB each, against 126 MB of L2)
    datas = [data.clone() for _ in range(reps + 1)]
    pars = [torch.zeros(n * 2, dtype=torch.uint8, device=dev) for _ in range(reps + 1)]
    enc_ms = timed(torch, lambda i: h.encode_batch_async(datas[i].data_ptr(), 1, 1, pars[i].data_ptr(), 2, n, stream),
                   reps + 1)
    par = pars[0]
    assert all(torch.equal(p_, par) for p_ in pars[1:])
    del datas
    # 15-bit codeword = (data & 31) << 10 | parity(10 bits, big-endian in 2 bytes); flip b in 0..4 distinct bits
    cwv = ((data.to(torch.int32) & 31) << 10) | (par.view(n, 2)[:, 0].to(torch.int32) << 8) | \
        par.view(n, 2)[:, 1].to(torch.int32)
    nb = torch.randint(0, 5, (n,), dtype=torch.int32, device=dev, generator=g)
    flips = torch.zeros(n, dtype=torch.int32, device=dev)
    for k in range(4):  # k-th flip: the r-th still-unset bit, r uniform in [0, 15 - k)
        r = torch.randint(0, 15 - k, (n,), dtype=torch.int32, device=dev, generator=g)
        seen = torch.zeros(n, dtype=torch.int32, device=dev)
        pick = torch.zeros(n, dtype=torch.int32, device=dev)
        for b in range(15):
            free = ((flips >> b) & 1) == 0
            pick |= torch.where(free & (seen == r), torch.full_like(pick, 1 << b), torch.zeros_like(pick))
            seen += free.to(torch.int32)
        flips |= torch.where(nb > k, pick, torch.zeros_like(pick))
    pop = torch.zeros(n, dtype=torch.int32, device=dev)
    for b in range(15):
        pop += (flips >> b) & 1
    rx = cwv ^ flips
    rdata0 = (((rx >> 10) & 31) | (data.to(torch.int32) & 0xE0)).to(torch.uint8)  # keep garbage in the top bits
    rpar = torch.stack([((rx >> 8) & 3).to(torch.uint8), (rx & 255).to(torch.uint8)], 1).contiguous()
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    cor = torch.empty(n, dtype=torch.int32, device=dev)
    rdatas = [rdata0.clone() for _ in range(reps + 1)]
    del pars
    rpars = [rpar.clone() for _ in range(reps + 1)]
    torch.cuda.synchronize()

    def dec(i):
        assert h.decode_batch_async(rdatas[i].data_ptr(), 1, 1, rpars[i].data_ptr(), 2, n, ok.data_ptr(),
                                    cor.data_ptr(), stream=stream)

    dec(reps)
    torch.cuda.synchronize()
    dec_ms = timed(torch, dec, reps)
    okb = ok.bool()
    le3 = pop <= 3
    assert bool(okb[le3].all().item()), "a BCH(15,5) word with <= 3 bit errors failed"
    assert bool((cor[le3] == pop[le3]).all().item()), "corrected_num differs from the injected bit count"
    assert bool((rdatas[0][le3] == (data[le3] & 31)).all().item()), "data byte not restored / pad bits not cleared"
    te, td = sum(enc_ms[1:]) / reps, sum(dec_ms) / reps
    emit(args, {"workload": "configs[3]: BCH(15,5) t=3, 0-4 flipped bits per codeword", "codewords": n,
                "encode_ms": te, "decode_ms": td, "encode_Gcw_per_s": n / te / 1e6, "decode_Gcw_per_s": n / td / 1e6,
                "encode_alg_GBps": n * 3 / te / 1e6, "decode_alg_GBps": n * (3 + 1 + 5) / td / 1e6,
                "encode_hbm_frac": n * 3 / te / 1e6 / hbm_peak(), "decode_hbm_frac": n * 9 / td / 1e6 / hbm_peak(),
                "ok_fraction": float(okb.float().mean().item()),
                "ok_fraction_of_4_bit_errors": float(okb[pop == 4].float().mean().item()),
                "mean_bit_errors": float(pop.float().mean().item())})
    h.close()


def run_bch31(args, torch, api, dev):
    """BCH(31,16) t=3 ("also BCH(31,k)"): n > 15: decoded through the remainder-indexed map (bch.cu)."""
    n = args.ncw or 64 * 1024 * 1024
    stream = torch.cuda.current_stream().cuda_stream
    h = api.Poporon.bch(5, 0x25, 3)
    g = torch.Generator(device=dev)
    g.manual_seed(0x5EED0431)
    data = torch.randint(0, 256, (n, 2), dtype=torch.uint8, device=dev, generator=g)  # 16 data bits, big-endian
    par = torch.zeros((n, 2), dtype=torch.uint8, device=dev)
    reps = args.reps
    enc_ms = timed(torch, lambda i: h.encode_batch_async(data.data_ptr(), 2, 2, par.data_ptr(), 2, n, stream), reps + 1)
    d32 = (data[:, 0].to(torch.int64) << 8) | data[:, 1].to(torch.int64)
    p32 = ((par[:, 0].to(torch.int64) << 8) | par[:, 1].to(torch.int64)) & 0x7FFF
    cw = (d32 << 15) | p32
    nb = torch.randint(0, 4, (n,), dtype=torch.int64, device=dev, generator=g)
    flips = torch.zeros(n, dtype=torch.int64, device=dev)
    for k in range(3):  # bits may coincide (then fewer than nb errors): the count below is the true weight
        b = torch.randint(0, 31, (n,), dtype=torch.int64, device=dev, generator=g)
        flips ^= torch.where(nb > k, torch.ones_like(b) << b, torch.zeros_like(b))
    pop = torch.zeros(n, dtype=torch.int64, device=dev)
    for b in range(31):
        pop += (flips >> b) & 1
    rx = cw ^ flips
    rd = torch.stack([((rx >> 23) & 255).to(torch.uint8), ((rx >> 15) & 255).to(torch.uint8)], 1).contiguous()
    rp = torch.stack([((rx >> 8) & 127).to(torch.uint8), (rx & 255).to(torch.uint8)], 1).contiguous()
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    cor = torch.empty(n, dtype=torch.int32, device=dev)
    rds = [rd.clone() for _ in range(reps + 1)]
    torch.cuda.synchronize()

    def dec(i):
        assert h.decode_batch_async(rds[i].data_ptr(), 2, 2, rp.data_ptr(), 2, n, ok.data_ptr(), cor.data_ptr(),
                                    stream=stream)

    dec(reps)
    torch.cuda.synchronize()
    dec_ms = timed(torch, dec, reps)
    assert bool(ok.bool().all().item()), "a BCH(31,16) word with <= 3 bit errors failed"
    assert bool((cor.to(torch.int64) == pop).all().item()), "corrected_num differs from the injected bit count"
    assert bool((rds[0] == data).all().item()), "data bytes not restored"
    te, td = sum(enc_ms[1:]) / reps, sum(dec_ms) / reps
    emit(args, {"workload": "BCH(31,16) t=3, 0-3 flipped bits per codeword (remainder-indexed decode map)", "codewords": n,
                "encode_ms": te, "decode_ms": td, "encode_Gcw_per_s": n / te / 1e6, "decode_Gcw_per_s": n / td / 1e6,
                "encode_alg_GBps": n * 4 / te / 1e6, "decode_alg_GBps": n * (4 + 2 + 5) / td / 1e6,
                "mean_bit_errors": float(pop.float().mean().item())})
    h.close()


def run_load(args, torch, api, dev):
    n = args.ncw or 128 * 1024 * 1024
    stream = torch.cuda.current_stream().cuda_stream
    h = api.Poporon.rs_default()
    clean = torch.empty(n * K_CW, dtype=torch.uint8, device=dev)
    api.synth_fill(0, clean.data_ptr(), K_CW, K_MSG, n, 0x5EED0001, 0, stream)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    assert h.encode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, n, stream)
    ev[1].record()
    torch.cuda.synchronize()
    enc_ms = ev[0].elapsed_time(ev[1])
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    cor = torch.empty(n, dtype=torch.int32, device=dev)
    work = torch.empty_like(clean)
    for p_ppm in (0, 10000, 50000):
        times = []
        for rep in range(args.reps + 1):
            work.copy_(clean)
            if p_ppm:
                api.synth_inject(0, work.data_ptr(), K_CW, K_CW, n, 0x5EED0500 + rep, 0, 1, 0, 0, p_ppm, stream)
            if rep == args.reps:
                idx = sample_index(torch, n, dev)
                rx_s = work.view(n, K_CW)[idx].cpu().numpy()
            h.profile(True)
            ev[0].record()
            assert h.decode_batch_async(work.data_ptr(), K_CW, K_MSG, work.data_ptr() + K_MSG, K_CW, n, ok.data_ptr(),
                                        cor.data_ptr(), stream=stream)
            ev[1].record()
            torch.cuda.synchronize()
            prof = h.profile_read()
            h.profile(False)
            if rep:
                times.append(ev[0].elapsed_time(ev[1]))
        mism, ork, kind = oracle_compare(rx_s, work.view(n, K_CW)[idx].cpu().numpy(), ok[idx].cpu().numpy(),
                                         cor[idx].cpu().numpy())
        assert mism == 0, f"decode differs from the CPU {kind} on {mism} of {idx.numel()} sampled codewords"
        okb = ok.bool()
        ok_frac = float(okb.float().mean().item())
        mean_cor = float(cor.float().mean().item())
        dirty = float((cor > 0).float().mean().item()) + (1 - ok_frac)
        # property: what was reported ok is a codeword again (and equals the original unless miscorrected, which
        # needs > 16 errors); failed rows are untouched = still not codewords
        ok2 = torch.empty_like(ok)
        cor2 = torch.empty_like(cor)
        assert h.decode_batch_async(work.data_ptr(), K_CW, K_MSG, work.data_ptr() + K_MSG, K_CW, n, ok2.data_ptr(),
                                    cor2.data_ptr(), stream=stream)
        torch.cuda.synchronize()
        assert bool((cor2[okb] == 0).all().item()) and bool(ok2[okb].all().item()), "a corrected row is not a codeword"
        same = torch.cat([(work.view(n, K_CW)[a:a + (1 << 24)] == clean.view(n, K_CW)[a:a + (1 << 24)]).all(1)
                          for a in range(0, n, 1 << 24)])
        restored_frac = float(same.float().mean().item())
        assert bool((~ok2.bool() == ~okb).all().item()), "second decode changed the verdict of a failed row"
        t = sum(times) / len(times)
        alg = n * (K_CW + 5 + mean_cor)
        emit(args, {"workload": "configs[4] single-GPU leg: RS(255,223) decode, i.i.d. symbol error rate %.0f %%"
                                % (p_ppm / 1e4), "codewords": n, "decode_ms": t, "payload_GBps": n * K_MSG / t / 1e6,
                    "alg_GBps": alg / t / 1e6, "hbm_frac": alg / t / 1e6 / hbm_peak(), "ok_fraction": ok_frac,
                    "restored_fraction": restored_frac, "mean_corrected": mean_cor, "dirty_fraction": dirty,
                    "encode_ms": enc_ms, "encode_payload_GBps": n * K_MSG / enc_ms / 1e6,
                    "verify": {"codewords": int(idx.numel()), "mismatches": mism, "against": kind,
                               "oracle_ok_fraction": ork, "fields": "data, parity, ok, corrected_num"},
                    "kernels": {k: {"launches": c, "avg_ms": m / c} for k, (c, m) in prof.items()}})
    h.close()


def run_load5(args, torch, api, dev):
    """BASELINE configs[4] as written: 128 Mi codewords in total, sharded contiguously over --gpus N devices (strong
    scaling), i.i.d. symbol-error rates 0 / 1 / 5 %, next to the reference's CPU path on all host cores.  One process:
    the devices are driven from one host thread through the asynchronous device-pointer API (no collective, no NCCL);
    the time of a pass is the largest per-device time between CUDA events on that device's stream."""
    import time

    import numpy as np

    N = args.gpus
    total = args.ncw or 128 * 1024 * 1024
    per = total // N
    hs, cleans, works, oks, cors, streams = [], [], [], [], [], []
    for d in range(N):
        torch.cuda.set_device(d)
        dv = torch.device("cuda", d)
        st = torch.cuda.current_stream(dv).cuda_stream
        h = api.Poporon.rs_default()
        clean = torch.empty(per * K_CW, dtype=torch.uint8, device=dv)
        api.synth_fill(d, clean.data_ptr(), K_CW, K_MSG, per, 0x5EED0001, d * per, st)
        assert h.encode_batch_async(clean.data_ptr(), K_CW, K_MSG, clean.data_ptr() + K_MSG, K_CW, per, st)
        hs.append(h), cleans.append(clean), works.append(torch.empty_like(clean)), streams.append(st)
        oks.append(torch.empty(per, dtype=torch.uint8, device=dv))
        cors.append(torch.empty(per, dtype=torch.int32, device=dv))
    for d in range(N):
        torch.cuda.synchronize(d)
    # CPU arm: the compiled reference (else the restatement) on all cores, on a bounded sample per load
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oraclelib import Oracle, have_reference

    kind = "reference" if have_reference() else "port"
    orc = Oracle(kind)
    cores = os.cpu_count() or 1
    sample = min(per, cores * 20000)
    for p_ppm in (0, 10000, 50000):
        times = []
        for rep in range(args.reps + 1):
            evs = []
            for d in range(N):
                torch.cuda.set_device(d)
                works[d].copy_(cleans[d])
                if p_ppm:
                    api.synth_inject(d, works[d].data_ptr(), K_CW, K_CW, per, 0x5EED0500 + rep, d * per, 1, 0, 0, p_ppm,
                                     streams[d])
            for d in range(N):
                torch.cuda.synchronize(d)
            if rep == args.reps:
                rx_s = works[0].view(per, K_CW)[:sample].cpu().numpy()
            for d in range(N):
                torch.cuda.set_device(d)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                assert hs[d].decode_batch_async(works[d].data_ptr(), K_CW, K_MSG, works[d].data_ptr() + K_MSG, K_CW, per,
                                                oks[d].data_ptr(), cors[d].data_ptr(), stream=streams[d])
                e1.record()
                evs.append((e0, e1))
            for d in range(N):
                torch.cuda.synchronize(d)
            if rep:
                times.append(max(e0.elapsed_time(e1) for e0, e1 in evs))
        ok_frac = float(np.mean([float(o.float().mean().item()) for o in oks]))
        mean_cor = float(np.mean([float(c.float().mean().item()) for c in cors]))
        t0 = time.perf_counter()
        wd, wp, wok, wcor = orc.rs_decode(np.ascontiguousarray(rx_s[:, :K_MSG]), np.ascontiguousarray(rx_s[:, K_MSG:]),
                                          threads=cores)
        cpu_s = time.perf_counter() - t0
        got = works[0].view(per, K_CW)[:sample].cpu().numpy()
        mism = int(((got[:, :K_MSG] != wd).any(1) | (got[:, K_MSG:] != wp).any(1) |
                    (oks[0][:sample].cpu().numpy() != wok) | (cors[0][:sample].cpu().numpy().astype(np.uint32) != wcor)).sum())
        assert mism == 0, f"decode differs from the CPU {kind} on {mism} sampled codewords"
        t = sum(times) / len(times)
        emit(args, {"workload": "configs[4]: RS(255,223) decode, 128 Mi codewords in total, i.i.d. symbol error rate %.0f %%"
                                % (p_ppm / 1e4), "gpus": N, "codewords_total": per * N, "codewords_per_gpu": per,
                    "scaling": "strong", "decode_ms": t, "payload_GBps": per * N * K_MSG / t / 1e6,
                    "alg_GBps": per * N * (K_CW + 5 + mean_cor) / t / 1e6,
                    "hbm_frac_per_gpu": per * (K_CW + 5 + mean_cor) / t / 1e6 / hbm_peak(), "ok_fraction": ok_frac,
                    "mean_corrected": mean_cor,
                    "cpu": {"kind": kind, "cores": cores, "sample": sample, "payload_GBps": sample * K_MSG / cpu_s / 1e9,
                            "mismatches_vs_gpu": mism}})
    for h in hs:
        h.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("which", nargs="*", default=["erasure", "bch", "load"])
    ap.add_argument("--gpus", type=int, default=1, help="load5: devices the 128 Mi codewords are sharded over")
    ap.add_argument("--ncw", type=int, default=0, help="codewords (0 = the configuration's own size)")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch

    import libpoporon_b200.api as api

    if not torch.cuda.is_available():
        raise SystemExit("bench_configs.py needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    for w in args.which:
        {"erasure": run_erasure, "bch": run_bch, "bch31": run_bch31, "load": run_load, "load5": run_load5}[w](args, torch, api, dev)
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
