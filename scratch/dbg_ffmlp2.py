import sys, os
sys.path.insert(0, 'tests'); sys.path.insert(0, 'torch-ngp_b200'); sys.path.insert(0, '.')
import numpy as np, torch
from oracle import cpu_ref
import test_gpu_ffmlp as T
cfg=(32,16,64,2)
in_dim,out_dim,hid,nl = cfg
for B in [256, 1024, 1000, 128*200]:
    torch.manual_seed(B+1)
    x=(torch.randn(B,in_dim,device='cuda')*0.7).half(); g=(torch.randn(B,out_dim,device='cuda')*0.5).half()
    w=T._weights(in_dim,out_dim,hid,nl,seed=2)
    w16=w.half().float().cpu().numpy().astype(np.float64)
    gx_ref,gw_ref=cpu_ref.ffmlp_backward(g.float().cpu().numpy(),x.float().cpu().numpy(),w16,in_dim,out_dim,hid,nl)
    for rep in range(3):
        gx,gw=T._bwd(g,x,w,in_dim,out_dim,hid,nl); torch.cuda.synchronize()
        gxn=gx.float().cpu().numpy(); gwn=gw.cpu().numpy()
        e=np.abs(gxn-gx_ref).max(1)/np.abs(gx_ref).max()
        nt=(B+127)//128
        pt=[float(e[t*128:(t+1)*128].max()) for t in range(min(nt,12))]
        bad=[int(i) for i in np.nonzero(e>2e-3)[0][:20]]
        print(B,rep,'gx per-tile err',np.round(pt,4),'bad rows',bad, 'gw err', np.abs(gwn-gw_ref).max()/np.abs(gw_ref).max())
