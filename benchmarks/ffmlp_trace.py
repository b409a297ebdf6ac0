"""Timeline of one CTA of the multi-pipeline FFMLP backward (developer tool, not a bench).

  python benchmarks/ffmlp_trace.py --build      # here (no GPU): ffmlp.cu with -DNGP_FFMLP_TRACE -> benchmarks/_build/libngp_trace.so
  python benchmarks/ffmlp_trace.py              # on the GPU box: runs one backward, prints CTA 0's event log in cycles

Events: issuer warps (slot 0 layer products, slot 1 weight gradients): 1xx = poll hit (xx = ready mask), 2pr = commit
issued for pipeline p round r.  Pipelines (slot 2+p): 300 tile start, 301 loads landed, 40l woke up for H_l, 50l H_l
written + arrive, 60l woke up for G_l, 70l G_l written + arrive, 800 woke up for dX, 900 tile done.
"""
import ctypes
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "torch-ngp_b200")
OUT = os.path.join(ROOT, "benchmarks", "_build")
LIB = os.path.join(OUT, "libngp_trace.so")


def build():
    sys.path.insert(0, PKG)
    import build as b
    b.build()
    os.makedirs(OUT, exist_ok=True)
    obj = os.path.join(OUT, "ffmlp_trace.o")
    subprocess.run([b.NVCC, "-c", os.path.join(b.CSRC, "ffmlp.cu"), "-o", obj, "-DNGP_FFMLP_TRACE"] + b.NVCC_FLAGS, check=True)
    objs = [o for o in glob.glob(os.path.join(b.OBJ, "*.o")) if not o.endswith("ffmlp.cu.o")] + [obj]
    subprocess.run([b.NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"], check=True)
    print(LIB)


def main():
    os.environ["NGP_B200_LIB"] = LIB
    sys.path.insert(0, ROOT)
    sys.path.insert(0, PKG)
    import numpy as np
    import torch
    import ffmlp
    import ngp_abi as abi
    be = ffmlp._backend
    lib = ctypes.CDLL(LIB)
    B, i, o, h, nl = 1 << 20, 32, 16, 64, 3
    x = (torch.randn(B, i, device="cuda") * 0.5).half()
    g = (torch.randn(B, o, device="cuda") * 0.1).half()
    nw = h * (i + h * (nl - 1) + o)
    w = (torch.rand(nw, device="cuda") * 2 - 1) * np.sqrt(3 / h)
    gx = torch.empty_like(x)
    gw = torch.empty(nw, device="cuda")
    for issuers in (3,):
        abi.set_option("ffmlp_bwd_issuers", issuers)
        if os.environ.get("NGP_ABLATE"):  # debug builds only (-DNGP_FFMLP_DEBUG)
            abi.set_option("ffmlp_ablate", int(os.environ["NGP_ABLATE"]))
        for _ in range(2):
            be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
        lib.ngp_ffmlp_trace_dump(None, None, 1)
        be.ffmlp_backward(g, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
        ev = np.zeros((8, 1024), dtype=np.uint64)
        cnt = np.zeros(8, dtype=np.uint32)
        lib.ngp_ffmlp_trace_dump(ev.ctypes.data_as(ctypes.c_void_p), cnt.ctypes.data_as(ctypes.c_void_p), 1)
        rows = []
        for s in range(8):
            for k in range(int(cnt[s])):
                rows.append((int(ev[s, k]) >> 12, s, int(ev[s, k]) & 0xfff))
        rows.sort()
        t0 = rows[0][0]
        print(f"==== issuers={issuers}: {len(rows)} events, CTA 0 ====")
        # middle of the run: tiles 6..8 of pipeline 0
        starts = [t for (t, s, tag) in rows if s == 2 and tag == 0x300]
        lo, hi = (starts[6], starts[8]) if len(starts) > 8 else (t0, t0 + 40000)
        names = {0: "issA", 1: "issB", 2: "p0", 3: "p1", 4: "p2", 5: "p3"}
        for (t, s, tag) in rows:
            if lo <= t <= hi:
                print(f"{t - lo:7d} {'  ' * 6 * s}{names.get(s, s)}:{tag:03x}")
        per_tile = np.diff(starts)
        print("pipeline 0 tile period (cycles): median", int(np.median(per_tile)), "min", int(per_tile.min()), "max", int(per_tile.max()))
        # how long each stage takes on pipeline 0, median over tiles
        p0 = [(t, tag) for (t, s, tag) in rows if s == 2]
        seg = {}
        for (ta, a), (tb, b2) in zip(p0[:-1], p0[1:]):
            seg.setdefault((a, b2), []).append(tb - ta)
        for k, v in seg.items():
            print(f"   p0 {k[0]:03x} -> {k[1]:03x}: median {int(np.median(v)):6d}  (n={len(v)})")


if __name__ == "__main__":
    build() if "--build" in sys.argv else main()
