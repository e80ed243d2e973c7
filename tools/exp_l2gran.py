#!/usr/bin/env python
"""Experiment: does cudaLimitMaxL2FetchGranularity change the stream kernels' time / DRAM traffic?
usage: tools/exp_l2gran.py GRAN   (32, 64 or 128); run under ncu --metrics dram__bytes_read.sum for the traffic."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import libpoporon_b200.api as api
gran = int(sys.argv[1])
torch.cuda.set_device(0)
torch.zeros(1, device="cuda")
rt = ctypes.CDLL("libcudart.so.12")
rc = rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran))
val = ctypes.c_size_t(0)
rt.cudaDeviceGetLimit(ctypes.byref(val), 5)
n = 4 * 1024 * 1024
h = api.Poporon.rs_default()
s = torch.cuda.current_stream().cuda_stream
buf = [torch.empty(n * 255, dtype=torch.uint8, device="cuda") for _ in range(3)]
for b in buf:
    api.synth_fill(0, b.data_ptr(), 255, 223, n, 1, 0, s)
ok = torch.empty(n, dtype=torch.uint8, device="cuda")
cor = torch.empty(n, dtype=torch.int32, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(7)]
for i in range(6):
    b = buf[i % 3]
    ev[i].record()
    if i < 3:
        h.encode_batch_async(b.data_ptr(), 255, 223, b.data_ptr() + 223, 255, n, s)
    else:
        h.decode_batch_async(b.data_ptr(), 255, 223, b.data_ptr() + 223, 255, n, ok.data_ptr(), cor.data_ptr(), stream=s)
ev[6].record()
torch.cuda.synchronize()
print("gran rc", rc, "limit", val.value, "encode ms", [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(3)],
      "clean decode ms", [round(ev[i].elapsed_time(ev[i + 1]), 3) for i in range(3, 6)])
