import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import numpy as np, torch, ffmlp, ngp_abi as abi
be = ffmlp._backend
def t(fn, iters=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3
for (B, i, o, h, nl) in [(589824, 32, 16, 64, 3), (589824, 32, 16, 64, 2)]:
    x = (torch.randn(B, i, device="cuda") * 0.5).half()
    g = (torch.randn(B, o, device="cuda") * 0.1).half()
    nw = h * (i + h * (nl - 1) + o)
    w = ((torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h))
    gx = torch.empty_like(x); gw = torch.empty(nw, device="cuda")
    y = torch.empty(B, o, device="cuda", dtype=torch.half)
    print(f"net nl={nl} fwd: {t(lambda: be.ffmlp_forward(x, w, B, i, o, h, nl, 0, 6, None, y)):.1f} us")
    for iss in (2, 3):
        for split in (0, 1, 2):
            abi.set_option("ffmlp_bwd_issuers", iss); abi.set_option("ffmlp_bwd_split", split)
            us = t(lambda: be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw))
            print(f"net nl={nl} issuers {iss} split {split}: {us:.1f} us", flush=True)
