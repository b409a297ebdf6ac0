"""Micro-benchmark of the grid encoder kernels at BASELINE config 1 (2^18 points, L=16 F=2 T=2^19).
Times every kernel variant through the C ABI and, with --ref, the reference's own extension
(oracle/_ref/_gridencoder.so) on the same tensors.  Development tool; bench.py is the contract."""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))


def time_fn(fn, iters=50, warmup=10, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts = np.array(ts)
    return float(np.median(ts)), float(ts.min())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=1 << 18)
    ap.add_argument("--ref", action="store_true")
    ap.add_argument("--cold", action="store_true", help="flush L2 (256 MiB memset) between iterations")
    ap.add_argument("--coherent", action="store_true", help="points ordered along rays instead of uniform random")
    args = ap.parse_args()
    import gridencoder
    import ngp_abi as abi
    be = gridencoder._backend
    dev = "cuda"
    enc = gridencoder.GridEncoder(desired_resolution=2048).to(dev)
    offsets = enc.offsets
    B, L, C, D, H, S = args.B, 16, 2, 3, 16, np.log2(enc.per_level_scale)
    torch.manual_seed(0)
    if args.coherent:
        n_rays = B // 64
        o = torch.rand(n_rays, 1, 3, device=dev)
        d = torch.nn.functional.normalize(torch.randn(n_rays, 1, 3, device=dev), dim=-1)
        t = torch.arange(64, device=dev).view(1, 64, 1) * (3.4641 / 1024)
        x = (o + d * t).clamp(0, 1).reshape(B, 3).contiguous()
    else:
        x = torch.rand(B, 3, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if args.cold else None
    res = {"B": B, "cold": args.cold, "coherent": args.coherent}
    for dtype, tag in ((torch.float32, "f32"), (torch.float16, "f16")):
        s = 4 if dtype == torch.float32 else 2
        emb = ((torch.rand_like(enc.embeddings) - 0.5) * 2).to(dtype)
        a_fwd = B * (4 * D + 8 * L * C * s + L * C * s)
        a_bwd = B * (L * C * s + 4 * D + 8 * L * C * s) + emb.numel() * s
        for layout, ltag in ((1, "BLC"),):
            out = torch.empty(L * B * C, device=dev, dtype=dtype)
            g = torch.randn(L * B * C, device=dev).to(dtype)
            ge = torch.zeros_like(emb)
            DEF = dict(hashgrid_fwd_variant=-1, hashgrid_bwd_variant=-1, hashgrid_v3_minb_fwd=-1, hashgrid_v3_minb_bwd=-1, hashgrid_v3_mix_fwd=0, hashgrid_v3_mix_bwd=1, hashgrid_bwd_red4=0, hashgrid_bwd_warp_agg=1)

            def setopts(**kw):
                for k_, v_ in {**DEF, **kw}.items():
                    abi.set_option(k_, v_)
            for nm, kw in (("lg4", dict(hashgrid_fwd_variant=2)), ("v3_mix0", {}), ("v3_mix1", dict(hashgrid_v3_mix_fwd=1)),
                           ("v3_mix0_occ4", dict(hashgrid_v3_minb_fwd=4)), ("v3_mix0_occ6", dict(hashgrid_v3_minb_fwd=6))):
                setopts(**kw)
                med, mn = time_fn(lambda: be.grid_encode_forward(x, emb, offsets, out, B, D, C, L, S, H, None, 0, False, 0, layout=layout), flush=flush)
                res[f"fwd_{tag}_{nm}"] = dict(us=med, us_min=mn, gbs=a_fwd / med / 1e3)
            for nm, kw in (("lg4", dict(hashgrid_bwd_variant=2)), ("v3_mix0_red4_0", dict(hashgrid_v3_mix_bwd=0, hashgrid_bwd_red4=0)),
                           ("v3_mix0_red4_1", dict(hashgrid_v3_mix_bwd=0, hashgrid_bwd_red4=1)), ("v3_mix1_red4_0", {}), ("v3_mix1_red4_1", dict(hashgrid_bwd_red4=1)),
                           ("v3_mix1_red4_0_occ6", dict(hashgrid_v3_minb_bwd=6)), ("v3_mix1_red4_0_noagg", dict(hashgrid_bwd_warp_agg=0))):
                setopts(**kw)

                def bwd():
                    ge.zero_()
                    be.grid_encode_backward(g, x, emb, offsets, ge, B, D, C, L, S, H, None, None, 0, False, 0, layout=layout)
                med, mn = time_fn(bwd, flush=flush)
                res[f"bwd_{tag}_{nm}"] = dict(us=med, us_min=mn, gbs=a_bwd / med / 1e3)
            setopts()
        if args.ref:
            from oracle import ref_ext
            ref = ref_ext.load("_gridencoder")
            if ref is not None:
                out = torch.empty(L, B, C, device=dev, dtype=dtype)
                g = torch.randn(L, B, C, device=dev).to(dtype)
                ge = torch.zeros_like(emb)
                med, mn = time_fn(lambda: ref.grid_encode_forward(x, emb, offsets, out, B, D, C, L, S, H, None, 0, False, 0), flush=flush)
                res[f"ref_fwd_{tag}_kernel"] = dict(us=med, us_min=mn, gbs=a_fwd / med / 1e3)
                med, mn = time_fn(lambda: (ref.grid_encode_forward(x, emb, offsets, out, B, D, C, L, S, H, None, 0, False, 0),
                                           out.permute(1, 0, 2).reshape(B, L * C)), flush=flush)
                res[f"ref_fwd_{tag}_with_permute"] = dict(us=med, us_min=mn, gbs=a_fwd / med / 1e3)

                def rbwd():
                    ge.zero_()
                    ref.grid_encode_backward(g, x, emb, offsets, ge, B, D, C, L, S, H, None, None, 0, False, 0)
                med, mn = time_fn(rbwd, flush=flush)
                res[f"ref_bwd_{tag}_kernel"] = dict(us=med, us_min=mn, gbs=a_bwd / med / 1e3)
        res[f"alg_bytes_{tag}"] = dict(fwd=a_fwd, bwd=a_bwd)
    for k, v in res.items():
        print(k, json.dumps(v))


if __name__ == "__main__":
    main()
