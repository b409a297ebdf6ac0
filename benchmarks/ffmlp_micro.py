"""Micro-benchmark of the tcgen05 FFMLP kernels through the C ABI vs the reference's own extension
(oracle/_ref/_ffmlp.so) on identical tensors.  Shapes: config 5 (SDF: 32->64x3->16, B=2^20), the NeRF sigma net
(32->64x2->16) and the reference's own micro-benchmark shape (testing/test_ffmlp.py:100-233: 16->64x2->16, B=2^21).
Development tool; bench.py is the contract."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))


def time_fn(fn, iters=30, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts)), float(np.min(ts))


def main():
    import ffmlp
    from oracle import ref_ext
    be = ffmlp._backend
    ref = ref_ext.load("_ffmlp") if "--ref" in sys.argv else None
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}
    for (B, i, o, h, nl) in [(1 << 20, 32, 16, 64, 3), (1 << 20, 32, 16, 64, 2), (1 << 21, 16, 16, 64, 2), (1 << 18, 32, 16, 64, 2), (1 << 20, 32, 16, 128, 2)]:
        x = (torch.randn(B, i, device="cuda") * 0.5).half()
        g = (torch.randn(B, o, device="cuda") * 0.1).half()
        nw = h * (i + h * (nl - 1) + o)
        w = ((torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h))
        w16 = w.half()
        y = torch.empty(B, o, device="cuda", dtype=torch.half)
        gx = torch.empty_like(x)
        gw = torch.empty(nw, device="cuda")
        flops_f = 2 * B * (i * h + (nl - 1) * h * h + h * o)
        t_f = time_fn(lambda: be.ffmlp_inference(x, w, B, i, o, h, nl, 0, 6, None, y))
        t_b = time_fn(lambda: be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, False, None, None, gw))
        t_bx = time_fn(lambda: be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw))
        res = {"B": B, "net": f"{i}->{h}x{nl}->{o}", "fwd_us": t_f[0], "fwd_tflops": flops_f / t_f[0] / 1e6, "fwd_frac_tensor": flops_f / t_f[0] / 1e6 / peaks["bf16_tflops"],
               "fwd_gbs": B * (i + o) * 2 / t_f[0] / 1e3, "bwd_us": t_b[0], "bwd_dx_us": t_bx[0],
               "bwd_dx_tflops_alg": (2 * flops_f) / t_bx[0] / 1e6, "bwd_dx_tflops_incl_recompute": (3 * flops_f) / t_bx[0] / 1e6}
        if ref is not None and h != 128:
            Bp = B
            ref.allocate_splitk(nl + 1)
            fb = torch.empty(nl, Bp, h, device="cuda", dtype=torch.half)
            bb = torch.zeros(nl, Bp, h, device="cuda", dtype=torch.half)
            gw16 = torch.zeros_like(w16)
            r_f = time_fn(lambda: ref.ffmlp_forward(x, w16, Bp, i, o, h, nl, 0, 6, fb, y))
            r_i = time_fn(lambda: ref.ffmlp_inference(x, w16, Bp, i, o, h, nl, 0, 6, fb[0], y))
            r_b = time_fn(lambda: ref.ffmlp_backward(g, x, w16, fb, Bp, i, o, h, nl, 0, 6, True, bb, gx, gw16))
            res.update({"ref_fwd_train_us": r_f[0], "ref_inference_us": r_i[0], "ref_bwd_dx_us": r_b[0]})
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
