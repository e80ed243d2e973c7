#!/usr/bin/env python
"""Per-kernel decode times of RS(255,223) / RS(255,239) / RS(255,247) at 0..t errors.  usage: PYTHONPATH=. tools/dec_probe.py [ncw]"""
import os
import sys
import torch
import libpoporon_b200.api as api

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4 << 20
dev = torch.device("cuda:0")
st = torch.cuda.current_stream().cuda_stream
for nr in (32, 16, 8):
    K = 255 - nr
    with api.Poporon.rs(8, 0x11D, 1, 1, nr) as h:
        rec = torch.empty(n * 255, dtype=torch.uint8, device=dev)
        api.synth_fill(0, rec.data_ptr(), 255, K, n, 1234, 0, st)
        assert h.encode_batch_async(rec.data_ptr(), 255, K, rec.data_ptr() + K, 255, n, stream=st)
        clean = rec.clone()
        ok = torch.empty(n, dtype=torch.uint8, device=dev)
        cor = torch.empty(n, dtype=torch.int32, device=dev)
        for rep in range(3):
            rec.copy_(clean)
            api.synth_inject(0, rec.data_ptr(), 255, 255, n, 99 + rep, 0, 0, 0, nr // 2, 0, st)
            if rep == 1:
                h.profile(True)
            assert h.decode_batch_async(rec.data_ptr(), 255, K, rec.data_ptr() + K, 255, n, ok.data_ptr(), cor.data_ptr(), stream=st)
            torch.cuda.synchronize()
        prof = h.profile_read()
        if not os.environ.get("DEC_PROBE_NOCHECK"):
            assert bool(ok.all().item()) and torch.equal(rec, clean)
        print(f"RS(255,{K})", {k: round(v[1] / v[0], 3) for k, v in prof.items()}, "mean errors", float(cor.float().mean().item()))
