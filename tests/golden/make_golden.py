"""Generates tests/golden/*.npz: small input/output vectors produced by the REFERENCE's own CUDA extensions
(lizuoyue/torch-ngp @ fb7ee66, compiled unmodified by oracle/build_ref_ext.py into oracle/_ref/) on a B200.

    gpurun -- python tests/golden/make_golden.py        (needs a GPU and oracle/_ref/*.so)

The fixtures pin the CPU oracle (tests/test_golden.py, no GPU needed) and, through it and directly, the new kernels
(tests/test_golden.py -m gpu).  Inputs are seeded; every array the reference call reads or writes is stored.
Ray-marching outputs are stored per ray in ray order (the reference places segments in atomicAdd arrival order,
raymarching.cu:405-406, so its raw buffers are not reproducible run to run).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "torch-ngp_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import scenes  # noqa: E402
from oracle import cpu_ref, ref_ext  # noqa: E402

DEV = "cuda"


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def save(name, **arrays):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **{k: (v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)) for k, v in arrays.items()})
    print(f"{name}.npz  {os.path.getsize(path) / 1024:.1f} KiB")


def grid():
    ref = ref_ext.load("_gridencoder")
    offsets, pls = cpu_ref.make_offsets(num_levels=8, log2_hashmap_size=12, base_resolution=8, desired_resolution=256)
    L, C, D, H, S = 8, 2, 3, 8, float(np.log2(pls))
    B = 384
    rng = np.random.default_rng(0)
    x = rng.random((B, D)).astype(np.float32)
    x[:4] = [[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [0.999, 0.001, 0.25]]
    emb = rng.standard_normal((int(offsets[-1]), C)).astype(np.float32)
    g = rng.standard_normal((L, B, C)).astype(np.float32)
    for dt, tag in ((torch.float32, "f32"), (torch.float16, "f16")):
        e = T(emb).to(dt)
        out = torch.empty(L, B, C, device=DEV, dtype=dt)
        dy = torch.empty(B, L * D * C, device=DEV, dtype=dt)
        ref.grid_encode_forward(T(x), e, T(offsets), out, B, D, C, L, S, H, dy, 0, False, 0)
        ge = torch.zeros_like(e)
        gi = torch.zeros(B, D, device=DEV, dtype=dt)
        ref.grid_encode_backward(T(g).to(dt), T(x), e, T(offsets), ge, B, D, C, L, S, H, dy, gi, 0, False, 0)
        torch.cuda.synchronize()
        # exp2f(level * S) as the GPU evaluates it (the per-level scale; libm differs in the last ulp) -- lets the CPU oracle
        # reproduce the reference's cell positions exactly
        lv2 = torch.exp2(torch.arange(L, device=DEV, dtype=torch.float32) * torch.tensor(S, dtype=torch.float32, device=DEV))
        save(f"grid_{tag}", level_exp2=lv2, x=x, embeddings=e.float(), offsets=offsets, per_level_scale=pls, base_resolution=H, grad=T(g).to(dt).float(),
             outputs=out.float(), dy_dx=dy.float(), grad_embeddings=ge.float(), grad_inputs=gi.float())


def grid_tv():
    """grad_total_variation (gridencoder.cu:503-642): `inputs` in the table dtype (:505), grad accumulated into.  fp32 only: the
    reference's half path is a no-op (its atomicAdd(at::Half*) is an empty stub, :21-26; the module always calls it in fp32)."""
    ref = ref_ext.load("_gridencoder")
    offsets, pls = cpu_ref.make_offsets(num_levels=8, log2_hashmap_size=12, base_resolution=8, desired_resolution=256)
    L, C, D, H, S = 8, 2, 3, 8, float(np.log2(pls))
    B, weight = 2048, 0.37
    rng = np.random.default_rng(4)
    x = rng.random((B, D)).astype(np.float32)
    x[:4] = [[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [-0.25, 0.5, 0.5]]
    emb = rng.standard_normal((int(offsets[-1]), C)).astype(np.float32)
    lv2 = torch.exp2(torch.arange(L, device=DEV, dtype=torch.float32) * torch.tensor(S, dtype=torch.float32, device=DEV))
    out = {}
    for dt, tag in ((torch.float32, "f32"),):
        e, xi = T(emb).to(dt), T(x).to(dt)
        g = torch.zeros_like(e)
        ref.grad_total_variation(xi, e, g, T(offsets), weight, B, D, C, L, S, H, 0, False)
        torch.cuda.synchronize()
        out.update({f"x_{tag}": xi.float(), f"embeddings_{tag}": e.float(), f"grad_{tag}": g.float()})
    save("grid_tv", level_exp2=lv2, offsets=offsets, per_level_scale=pls, base_resolution=H, weight=weight, **out)


def raymarching():
    ref = ref_ext.load("_raymarching")
    H, C, bound = 32, 2, 2.0
    gridv = scenes.density_grid(H=H, cascade=C, bound=bound)
    o, d = scenes.camera_rays(192, seed=5, radius=1.6 * bound)
    aabb = np.array([-bound] * 3 + [bound] * 3, np.float32)
    N = o.shape[0]
    nears, fars = torch.empty(N, device=DEV), torch.empty(N, device=DEV)
    ref.near_far_from_aabb(T(o), T(d), T(aabb), N, 0.2, nears, fars)
    sph = torch.empty(N, 2, device=DEV)
    ref.sph_from_ray(T(o), T(d), 8.0, N, sph)
    rng = np.random.default_rng(1)
    coords = rng.integers(0, 1024, (512, 3)).astype(np.int32)
    mort = torch.empty(512, dtype=torch.int32, device=DEV)
    ref.morton3D(T(coords), 512, mort)
    inv = torch.empty(512, 3, dtype=torch.int32, device=DEV)
    ref.morton3D_invert(mort, 512, inv)
    bits = torch.empty(gridv.size // 8, dtype=torch.uint8, device=DEV)
    ref.packbits(T(gridv), gridv.size // 8, 0.01, bits)
    save("raymarching_utils", rays_o=o, rays_d=d, aabb=aabb, min_near=0.2, nears=nears, fars=fars, radius=8.0, sph=sph, coords=coords,
         morton=mort, coords_back=inv, grid=gridv, thresh=0.01, bitfield=bits, H=H, C=C, bound=bound)

    noises = rng.random(N).astype(np.float32)
    for dt_gamma, tag in ((0.0, "g0"), (1.0 / 128, "g128")):
        M = N * 256
        xyzs, dirs, deltas = torch.zeros(M, 3, device=DEV), torch.zeros(M, 3, device=DEV), torch.zeros(M, 2, device=DEV)
        rays = torch.empty(N, 3, dtype=torch.int32, device=DEV)
        counter = torch.zeros(2, dtype=torch.int32, device=DEV)
        ref.march_rays_train(T(o), T(d), bits, bound, dt_gamma, 256, N, C, H, M, nears, fars, xyzs, dirs, deltas, rays, counter, T(noises))
        torch.cuda.synchronize()
        r = rays.cpu().numpy()
        order = np.argsort(r[:, 0], kind="stable")
        r = r[order]
        cnt = r[:, 2].astype(np.int64)
        idx = np.repeat(r[:, 1].astype(np.int64), cnt) + (np.arange(cnt.sum()) - np.repeat(np.cumsum(cnt) - cnt, cnt))
        x_ro, d_ro, l_ro = xyzs.cpu().numpy()[idx], dirs.cpu().numpy()[idx], deltas.cpu().numpy()[idx]
        # composite over the ray-ordered samples with analytic sigma / colour
        off = (np.cumsum(cnt) - cnt).astype(np.int32)
        rays_ro = np.stack([np.arange(N, dtype=np.int32), off, cnt.astype(np.int32)], -1)
        Mr = int(cnt.sum())
        sig = (scenes.density(x_ro / bound) * rng.random(Mr)).astype(np.float32)
        rgb = scenes.colour(x_ro)
        ws, dep, img = torch.empty(N, device=DEV), torch.empty(N, device=DEV), torch.empty(N, 3, device=DEV)
        ref.composite_rays_train_forward(T(sig), T(rgb), T(l_ro), T(rays_ro), Mr, N, 1e-4, ws, dep, img)
        g_ws, g_dep, g_img = rng.standard_normal(N).astype(np.float32), rng.standard_normal(N).astype(np.float32), rng.standard_normal((N, 3)).astype(np.float32)
        gs, gc = torch.zeros(Mr, device=DEV), torch.zeros(Mr, 3, device=DEV)
        ref.composite_rays_train_backward(T(g_ws), T(g_dep), T(g_img), T(sig), T(rgb), T(l_ro), T(rays_ro), ws, dep, img, Mr, N, 1e-4, gs, gc)
        torch.cuda.synchronize()
        save(f"march_train_{tag}", rays_o=o, rays_d=d, bitfield=bits, bound=bound, dt_gamma=dt_gamma, max_steps=256, C=C, H=H, nears=nears, fars=fars,
             noises=noises, counter=counter, counts=cnt, xyzs=x_ro, dirs=d_ro, deltas=l_ro, rays=rays_ro, sigmas=sig, rgbs=rgb, T_thresh=1e-4,
             weights_sum=ws, depth=dep, image=img, grad_weights_sum=g_ws, grad_depth=g_dep, grad_image=g_img, grad_sigmas=gs, grad_rgbs=gc)

    # two iterations of the inference pair
    alive = torch.arange(N, dtype=torch.int32, device=DEV)
    rays_t = nears.clone()
    ws, dep, img = torch.zeros(N, device=DEV), torch.zeros(N, device=DEV), torch.zeros(N, 3, device=DEV)
    rec = {}
    for it in range(2):
        n_alive = alive.shape[0]
        n_step = 4
        Mi = n_alive * n_step
        xyzs, dirs, deltas = torch.zeros(Mi, 3, device=DEV), torch.zeros(Mi, 3, device=DEV), torch.zeros(Mi, 2, device=DEV)
        rec[f"alive_in{it}"], rec[f"t_in{it}"] = alive.clone(), rays_t.clone()
        ref.march_rays(n_alive, n_step, alive, rays_t, T(o), T(d), bound, 1.0 / 256, 256, C, H, bits, nears, fars, xyzs, dirs, deltas,
                       torch.zeros(n_alive, device=DEV))
        sig = T((scenes.density(xyzs.cpu().numpy() / bound)).astype(np.float32))
        rgb = T(scenes.colour(xyzs.cpu().numpy()))
        ref.composite_rays(n_alive, n_step, 1e-4, alive, rays_t, sig, rgb, deltas, ws, dep, img)
        torch.cuda.synchronize()
        rec.update({f"xyzs{it}": xyzs, f"dirs{it}": dirs, f"deltas{it}": deltas, f"sigmas{it}": sig, f"rgbs{it}": rgb, f"alive_out{it}": alive.clone(),
                    f"t_out{it}": rays_t.clone(), f"ws{it}": ws.clone(), f"depth{it}": dep.clone(), f"image{it}": img.clone()})
        alive = alive[alive >= 0].contiguous()
    save("march_infer", rays_o=o, rays_d=d, bitfield=bits, bound=bound, dt_gamma=1.0 / 256, max_steps=256, C=C, H=H, nears=nears, fars=fars,
         n_step=4, T_thresh=1e-4, **rec)


def ffmlp():
    ref = ref_ext.load("_ffmlp")
    rng = np.random.default_rng(2)
    for (i, o, h, nl) in ((32, 16, 64, 2), (32, 16, 64, 3)):
        B = 256
        x = (rng.standard_normal((B, i)) * 0.5).astype(np.float16)
        g = (rng.standard_normal((B, o)) * 0.1).astype(np.float16)
        nw = h * (i + h * (nl - 1) + o)
        w = ((rng.random(nw) * 2 - 1) * np.sqrt(3 / h)).astype(np.float16)
        ref.allocate_splitk(nl + 1)
        fb = torch.empty(nl, B, h, device=DEV, dtype=torch.half)
        y = torch.empty(B, o, device=DEV, dtype=torch.half)
        ref.ffmlp_forward(T(x), T(w), B, i, o, h, nl, 0, 6, fb, y)
        y_inf = torch.empty(B, o, device=DEV, dtype=torch.half)
        ref.ffmlp_inference(T(x), T(w), B, i, o, h, nl, 0, 6, fb[0].clone(), y_inf)
        bb = torch.zeros(nl, B, h, device=DEV, dtype=torch.half)
        gx = torch.zeros(B, i, device=DEV, dtype=torch.half)
        gw = torch.zeros(nw, device=DEV, dtype=torch.half)
        ref.ffmlp_backward(T(g), T(x), T(w), fb, B, i, o, h, nl, 0, 6, True, bb, gx, gw)
        torch.cuda.synchronize()
        save(f"ffmlp_{i}_{h}x{nl}_{o}", inputs=x.astype(np.float32), weights=w.astype(np.float32), grad=g.astype(np.float32), dims=np.array([i, o, h, nl]),
             outputs=y.float(), outputs_inference=y_inf.float(), forward_buffer=fb.float(), grad_inputs=gx.float(), grad_weights=gw.float())


def shenc():
    ref = ref_ext.load("_shencoder")
    rng = np.random.default_rng(3)
    d = rng.standard_normal((256, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    out = torch.empty(256, 16, device=DEV)
    ref.sh_encode_forward(T(d), out, 256, 3, 4, None)
    torch.cuda.synchronize()
    save("shencoder_deg4", dirs=d, outputs=out)
    # degree 8 (all 64 coefficients) with the reference's dy_dx table, on unit directions and on unnormalised points
    d8 = np.concatenate([d, rng.standard_normal((256, 3)).astype(np.float32) * 0.7])
    out8 = torch.empty(512, 64, device=DEV)
    dy_dx = torch.empty(512, 64 * 3, device=DEV)
    ref.sh_encode_forward(T(d8), out8, 512, 3, 8, dy_dx)
    torch.cuda.synchronize()
    save("shencoder_deg8", dirs=d8, outputs=out8, dy_dx=dy_dx)


if __name__ == "__main__":
    cpu_ref.build()
    only = sys.argv[1:]
    if only:  # e.g. `make_golden.py grid_tv`: (re)generate the named fixtures only
        for name in only:
            globals()[name]()
        sys.exit(0)
    grid()
    grid_tv()
    raymarching()
    ffmlp()
    shenc()
