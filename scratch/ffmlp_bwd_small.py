import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import numpy as np, torch, ffmlp, ngp_abi as abi
be = ffmlp._backend
for (B, i, o, h, nl) in [(1 << 16, 32, 16, 64, 3), (1 << 16, 32, 16, 64, 2)]:
    x = (torch.randn(B, i, device="cuda") * 0.5).half()
    g = (torch.randn(B, o, device="cuda") * 0.1).half()
    nw = h * (i + h * (nl - 1) + o)
    w = ((torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h))
    gx = torch.empty_like(x); gw = torch.empty(nw, device="cuda")
    abi.set_option("ffmlp_bwd_variant", 1)
    for split in (1, 2):
        abi.set_option("ffmlp_bwd_split", split)
        be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
        torch.cuda.synchronize()
        print("ok", nl, split, float(gw.abs().sum()))
