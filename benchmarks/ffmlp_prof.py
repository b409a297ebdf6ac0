"""One FFMLP backward (and forward) launch per net at the training batch, for `ncu -k regex:k_ffmlp -s <skip> -c <n>` captures.
Development tool (profiles/ holds the summaries)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import numpy as np  # noqa: E402


def main():
    import ffmlp
    be = ffmlp._backend
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 600000
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    for (i, o, h, nl) in [(32, 16, 64, 3), (32, 16, 64, 2)]:
        x = (torch.randn(B, i, device="cuda") * 0.5).half()
        g = (torch.randn(B, o, device="cuda") * 0.1).half()
        nw = h * (i + h * (nl - 1) + o)
        w = ((torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h))
        y = torch.empty(B, o, device="cuda", dtype=torch.half)
        gx = torch.empty_like(x)
        gw = torch.empty(nw, device="cuda")
        for _ in range(reps):
            be.ffmlp_forward(x, w, B, i, o, h, nl, 0, 6, None, y)
            be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
        torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
