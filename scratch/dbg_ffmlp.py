import sys, os
sys.path.insert(0, 'tests'); sys.path.insert(0, 'torch-ngp_b200'); sys.path.insert(0, '.')
import numpy as np, torch
from oracle import cpu_ref
import test_gpu_ffmlp as T
for cfg, B in [((32,16,64,2),1000), ((32,16,64,2),128), ((16,16,16,2),128)]:
    in_dim,out_dim,hid,nl = cfg
    torch.manual_seed(B+1)
    x=(torch.randn(B,in_dim,device='cuda')*0.7).half(); g=(torch.randn(B,out_dim,device='cuda')*0.5).half()
    w=T._weights(in_dim,out_dim,hid,nl,seed=2)
    gx,gw=T._bwd(g,x,w,in_dim,out_dim,hid,nl); torch.cuda.synchronize()
    w16=w.half().float().cpu().numpy().astype(np.float64)
    gx_ref,gw_ref=cpu_ref.ffmlp_backward(g.float().cpu().numpy(),x.float().cpu().numpy(),w16,in_dim,out_dim,hid,nl)
    gxn=gx.float().cpu().numpy(); gwn=gw.cpu().numpy()
    print(cfg,B,'gx err',np.abs(gxn-gx_ref).max()/np.abs(gx_ref).max(), 'nan', np.isnan(gxn).sum())
    e=np.abs(gxn-gx_ref); r,c=np.unravel_index(e.argmax(),e.shape); print(' worst row/col',r,c,gxn[r,c],gx_ref[r,c])
    print(' per-row max err (first 16 rows)', e.max(1)[:16].round(3))
    print(' per-col max err', e.max(0).round(3))
    o=0
    for i,(rows,K) in enumerate([(hid,in_dim)]+[(hid,hid)]*(nl-1)+[(out_dim,hid)]):
        a=gwn[o:o+rows*K].reshape(rows,K); b=gw_ref[o:o+rows*K].reshape(rows,K); o+=rows*K
        ee=np.abs(a-b); print(' dW',i,'err',ee.max()/np.abs(b).max(), 'rows', ee.max(1).round(2)[:8], 'cols', ee.max(0).round(2)[:8])
