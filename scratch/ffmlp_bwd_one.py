import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import numpy as np, torch, ffmlp
be = ffmlp._backend
B, i, o, h, nl = 1 << 20, 32, 16, 64, 3
x = (torch.randn(B, i, device="cuda") * 0.5).half()
g = (torch.randn(B, o, device="cuda") * 0.1).half()
nw = h * (i + h * (nl - 1) + o)
w = ((torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h))
gx = torch.empty_like(x); gw = torch.empty(nw, device="cuda")
y = torch.empty(B, o, device="cuda", dtype=torch.half)
for _ in range(3):
    be.ffmlp_inference(x, w, B, i, o, h, nl, 0, 6, None, y)
    be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
torch.cuda.synchronize()
print("ok")
