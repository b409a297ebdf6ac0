"""ncu target: one fwd + bwd of the hash grid for (f32, random points) and (f16, ray-ordered points)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import gridencoder, ngp_abi as abi
be = gridencoder._backend
dev = "cuda"
enc = gridencoder.GridEncoder(desired_resolution=2048).to(dev)
L, C, D, H, S = 16, 2, 3, 16, np.log2(enc.per_level_scale)
torch.manual_seed(0)
for coherent, dtype, B in ((False, torch.float32, 1 << 18), (True, torch.float16, 589824)):
    if coherent:
        n_rays = B // 64
        o = torch.rand(n_rays, 1, 3, device=dev)
        d = torch.nn.functional.normalize(torch.randn(n_rays, 1, 3, device=dev), dim=-1)
        t = torch.arange(64, device=dev).view(1, 64, 1) * (3.4641 / 1024)
        x = (o + d * t).clamp(0, 1).reshape(B, 3).contiguous()
    else:
        x = torch.rand(B, 3, device=dev)
    emb = ((torch.rand_like(enc.embeddings) - 0.5) * 2).to(dtype)
    out = torch.empty(L * B * C, device=dev, dtype=dtype)
    g = torch.randn(L * B * C, device=dev).to(dtype)
    ge = torch.zeros_like(emb)
    for _ in range(2):
        be.grid_encode_forward(x, emb, enc.offsets, out, B, D, C, L, S, H, None, 0, False, 0, layout=1)
        be.grid_encode_backward(g, x, emb, enc.offsets, ge, B, D, C, L, S, H, None, None, 0, False, 0, layout=1)
    torch.cuda.synchronize()
