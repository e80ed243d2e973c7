"""Timing probe of the general-parameter RS kernels (RS(255,239), RS(255,247)) on cuda:0: PYTHONPATH=. python tools/gen_probe.py"""
import time, torch, numpy as np
import libpoporon_b200.api as api
torch.cuda.set_device(0); dev=torch.device('cuda',0)
stream=torch.cuda.current_stream().cuda_stream
for (nr, n) in ((16, 4*1024*1024), (8, 4*1024*1024)):
    K=255-nr
    h=api.Poporon.rs(8,0x11D,1,1,nr)
    g=torch.Generator(device=dev); g.manual_seed(5)
    rec=torch.randint(0,256,(n,255),dtype=torch.uint8,device=dev,generator=g)
    ok=torch.zeros(n,dtype=torch.uint8,device=dev); cor=torch.zeros(n,dtype=torch.int32,device=dev)
    def enc(): assert h.encode_batch_async(rec.data_ptr(),255,K,rec.data_ptr()+K,255,n,stream)
    enc(); torch.cuda.synchronize()
    t0=time.time(); enc(); torch.cuda.synchronize(); te=(time.time()-t0)*1e3
    clean=rec.clone()
    # inject e ~ U{0..nr/2} errors
    e=torch.randint(0,nr//2+1,(n,),device=dev,generator=g)
    for k in range(nr//2):
        pos=torch.randint(0,255,(n,),device=dev,generator=g)
        mag=torch.randint(1,256,(n,),dtype=torch.int32,device=dev,generator=g).to(torch.uint8)
        m=(e>k)
        idx=torch.nonzero(m).squeeze(1)
        rec[idx,pos[idx]] ^= mag[idx]
    def dec(): assert h.decode_batch_async(rec.data_ptr(),255,K,rec.data_ptr()+K,255,n,ok.data_ptr(),cor.data_ptr(),stream=stream)
    t0=time.time(); dec(); torch.cuda.synchronize(); td=(time.time()-t0)*1e3
    print(f"RS(255,{K}) n={n}: encode {te:.1f} ms, decode {td:.1f} ms, ok {ok.float().mean().item():.4f}, restored {bool((rec==clean).all().item())}")
    h.close()
