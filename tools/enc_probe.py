#!/usr/bin/env python
"""Encoder kernel times by parity layout (packed records / split aligned / split odd).  usage: PYTHONPATH=. tools/enc_probe.py [ncw]"""
import sys
import torch
import libpoporon_b200.api as api

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4 << 20
dev = torch.device("cuda:0")
with api.Poporon.rs_default() as h:
    for name, dstride, poff, pstride, sep in (("packed", 255, 223, 255, False), ("split32", 223, 0, 32, True),
                                              ("split33", 223, 0, 33, True), ("split96+1", 223, 1, 96, True)):
        d = torch.randint(0, 256, (n * dstride + 64,), dtype=torch.uint8, device=dev)
        p = torch.zeros((n * pstride + 64,), dtype=torch.uint8, device=dev) if sep else d
        st = torch.cuda.current_stream().cuda_stream
        for rep in range(3):
            if rep == 1:
                h.profile(True)
            assert h.encode_batch_async(d.data_ptr(), dstride, 223, p.data_ptr() + poff, pstride, n, stream=st)
            torch.cuda.synchronize()
        prof = h.profile_read()
        print(name, {k: round(v[1] / v[0], 3) for k, v in prof.items()})
