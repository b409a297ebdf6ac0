"""In-job A/B of the FFMLP launch knobs (ngp_set_option) at the training batch.  Development tool."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import numpy as np, torch
import ffmlp, ngp_abi as abi
be = ffmlp._backend

def t(fn, it=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(it):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts))

B = 600000
for (i, o, h, nl) in [(32, 16, 64, 3), (32, 16, 64, 2)]:
    x = (torch.randn(B, i, device="cuda") * 0.5).half(); g = (torch.randn(B, o, device="cuda") * 0.1).half()
    nw = h * (i + h * (nl - 1) + o); w = ((torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h))
    y = torch.empty(B, o, device="cuda", dtype=torch.half); gx = torch.empty_like(x); gw = torch.empty(nw, device="cuda")
    fwd = lambda: be.ffmlp_inference(x, w, B, i, o, h, nl, 0, 6, None, y)
    bwd = lambda: be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
    print(f"net {i}->{h}x{nl}->{o}: default fwd {t(fwd):.1f} bwd {t(bwd):.1f}")
    for p in (3, 4, 5, 6, 7, 8):
        abi.set_option("ffmlp_fwd_pipes", p); print(f"   fwd pipes {p}: {t(fwd):.1f}")
    abi.set_option("ffmlp_fwd_pipes", 0)
    for ns in (1, 2):
        for iss in (2, 3):
            abi.set_option("ffmlp_bwd_split", ns); abi.set_option("ffmlp_bwd_issuers", iss)
            print(f"   bwd split {ns} issuers {iss}: {t(bwd):.1f}")
    abi.set_option("ffmlp_bwd_split", 0); abi.set_option("ffmlp_bwd_issuers", 3)
