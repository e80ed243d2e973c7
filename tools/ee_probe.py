import numpy as np, sys
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import libpoporon_b200.api as api
from oraclelib import Oracle, MODE_ERASURE, RS_DEFAULT
rng=np.random.default_rng(5); n=65536; size=223; F=10
orc=Oracle("port")
with api.Poporon.rs_default() as h:
    msg=rng.integers(0,256,(n,size),dtype=np.uint8); par=np.zeros((n,32),np.uint8)
    assert h.encode_batch(msg,size,size,par,32,n)
    cw=np.concatenate([msg,par],1)
    pos=np.zeros((n,32),np.uint32); cnt=np.full(n,F,np.uint32)
    pick=np.sort(np.argsort(rng.random((n,size)),axis=1)[:,:F],axis=1); pos[:,:F]=pick
    rows=np.repeat(np.arange(n),F).reshape(n,F); cw[rows,pick]^=rng.integers(0,256,(n,F),dtype=np.uint8)
    cw[:,250]^=0x5A  # one undeclared error in every row
    d=np.ascontiguousarray(cw[:,:size]); p=np.ascontiguousarray(cw[:,size:]); ok=np.zeros(n,np.uint8); cor=np.zeros(n,np.uint32)
    h.profile(True)
    assert h.decode_batch_erasure(d,size,size,p,32,n,pos,32,cnt,ok,cor)
    print({k:(v[0],round(v[1],3)) for k,v in h.profile_read().items()})
    wd,wp,wok,wcor=orc.rs_decode(cw[:,:size],cw[:,size:],RS_DEFAULT,MODE_ERASURE,pos,cnt,threads=8)
    print((ok==wok).all(),(cor==wcor).all(),(d==wd).all(),(p==wp).all(), wok.mean())
