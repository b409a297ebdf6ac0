"""The reference's OWN Python wrappers running on this repository's backend shims (SURVEY 8c plan (ii)).

oracle/build_ref_ext.py stages lizuoyue/torch-ngp's `gridencoder/grid.py`, `raymarching/raymarching.py` and `ffmlp/ffmlp.py`
byte for byte into oracle/_ref/py/ (build output, git-ignored, travels to the GPU box like the compiled extensions);
oracle/ref_ext.load_wrapper imports them with `_backend` = torch-ngp_b200/<pkg>/backend.py in place of the pybind11 module
(`import _gridencoder as _backend`, grid.py:22-25; MinkowskiEngine / matplotlib / turtle imports stubbed).  This is the one place
where the ops are called exactly the way the reference calls them: positional arguments only, [L, B, C] layout with the
transposes of grid.py:69,87, forward_buffer / backward_buffer / a dummy 1-element grad_inputs, half weights and half
grad_weights under autocast (custom_fwd(cast_inputs=torch.half)), batch padded to 128 rows.  Results are compared with this
repository's own wrapper modules on the same tensors.
"""
import warnings

import numpy as np
import pytest
import torch

import scenes

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _wrapper(pkg):
    from oracle import ref_ext
    import importlib
    backend = importlib.import_module(pkg)._backend
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # torch.cuda.amp.custom_fwd deprecation notices of the reference's decorators
        m = ref_ext.load_wrapper(pkg, backend)
    if m is None:
        pytest.skip("oracle/_ref/py not staged (python oracle/build_ref_ext.py, needs /root/reference)")
    return m


def test_reference_grid_encoder_module_on_shim():
    import gridencoder
    ref = _wrapper("gridencoder")
    kw = dict(num_levels=16, level_dim=2, base_resolution=16, log2_hashmap_size=19, desired_resolution=2048)
    torch.manual_seed(0)
    a = ref.GridEncoder(**kw).to(DEV)   # the reference's class (grid.py:754-843) ...
    b = gridencoder.GridEncoder(**kw).to(DEV)  # ... and this repository's
    assert torch.equal(a.offsets, b.offsets) and a.embeddings.shape == b.embeddings.shape and a.output_dim == b.output_dim
    with torch.no_grad():
        a.embeddings.uniform_(-1, 1)
        b.embeddings.copy_(a.embeddings)
    x = torch.rand(3, 7001, 3, device=DEV) * 2.2 - 1.1  # prefix dims, a few points outside [-1, 1]
    for autocast in (False, True):
        for req in (False, True):
            xa, xb = x.clone().requires_grad_(req), x.clone().requires_grad_(req)
            a.embeddings.grad = b.embeddings.grad = None
            with torch.autocast("cuda", dtype=torch.half, enabled=autocast), warnings.catch_warnings():
                warnings.simplefilter("ignore")
                ya, yb = a(xa, bound=1), b(xb, bound=1)
            assert ya.dtype == yb.dtype == (torch.half if autocast else torch.float32) and ya.shape == yb.shape == (3, 7001, 32)
            assert torch.equal(ya, yb)  # [L,B,C] + permute (grid.py:69) vs point-major kernels: the same numbers
            g = torch.randn_like(ya.float()).to(ya.dtype)
            ya.backward(g); yb.backward(g)
            tol = 1e-2 if autocast else 1e-5  # half table gradients are summed with half atomics (order-dependent)
            scale = b.embeddings.grad.abs().max().item()
            assert (a.embeddings.grad - b.embeddings.grad).abs().max().item() <= tol * scale
            if req:  # grid.py:91-99; the default input gradient IS the reference's (gridencoder.cu:143)
                assert torch.allclose(xa.grad, xb.grad, rtol=1e-3 if autocast else 1e-5, atol=(2e-2 if autocast else 1e-5) * xb.grad.abs().max().item())
                assert float(xa.grad[..., 1:].abs().max()) == 0.0 and float(xa.grad[..., 0].abs().max()) > 0.0
    # TV regulariser through the reference's method (grid.py:822-843)
    a.embeddings.grad = torch.zeros_like(a.embeddings); b.embeddings.grad = torch.zeros_like(b.embeddings)
    pts = torch.rand(5000, 3, device=DEV) * 2 - 1
    a.grad_total_variation(1e-3, pts, bound=1); b.grad_total_variation(1e-3, pts, bound=1)
    assert a.embeddings.grad.abs().max() > 0
    assert torch.allclose(a.embeddings.grad, b.embeddings.grad, rtol=1e-4, atol=1e-6 * a.embeddings.grad.abs().max().item())


@pytest.mark.parametrize("cfg", [(32, 16, 64, 2), (32, 3, 64, 3), (16, 1, 32, 2), (32, 3, 256, 2)])
def test_reference_ffmlp_module_on_shim(cfg):
    import ffmlp
    ref = _wrapper("ffmlp")
    in_dim, out_dim, hid, nl = cfg
    a = ref.FFMLP(in_dim, out_dim, hid, nl).to(DEV)     # ffmlp.py:99-168: seeds itself with manual_seed(42)
    b = ffmlp.FFMLP(in_dim, out_dim, hid, nl).to(DEV)
    assert torch.equal(a.weights, b.weights) and a.num_parameters == b.num_parameters
    torch.manual_seed(3)
    for B in (1000, 1024, 77):  # ragged: the reference pads to a multiple of 128 itself (ffmlp.py:156-159), and appends a whole tile when B % 128 == 0
        x = (torch.randn(B, in_dim, device=DEV) * 0.5)
        for req in (False, True):
            xa, xb = x.clone().requires_grad_(req), x.clone().requires_grad_(req)
            a.weights.grad = b.weights.grad = None
            a.train(); b.train()
            with torch.autocast("cuda", dtype=torch.half), warnings.catch_warnings():
                warnings.simplefilter("ignore")
                ya, yb = a(xa), b(xb)
            assert ya.shape == yb.shape == (B, out_dim) and ya.dtype == yb.dtype == torch.half
            assert torch.equal(ya, yb)
            g = torch.randn(B, out_dim, device=DEV).half() * 0.1
            ya.backward(g); yb.backward(g)
            # the reference's call hands over HALF weights and gets HALF grad_weights back (cast to the fp32 Parameter by
            # autograd); this repository's module asks for fp32: equal up to that one rounding
            scale = b.weights.grad.abs().max().item()
            assert (a.weights.grad - b.weights.grad).abs().max().item() <= 1e-3 * scale
            if req:
                assert torch.equal(xa.grad, xb.grad)
        a.eval(); b.eval()
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.half), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            assert torch.equal(a(x), b(x)) and torch.equal(a(x), yb.detach())  # ffmlp_inference == ffmlp_forward


def test_reference_raymarching_functions_on_shim():
    import raymarching as mine
    ref = _wrapper("raymarching")
    H, C, bound = 128, 2, 2.0
    grid = scenes.density_grid(H=H, cascade=C, bound=bound)
    o, d = scenes.camera_rays(4096, seed=21, radius=1.6 * bound)
    ro, rd = torch.from_numpy(o).to(DEV), torch.from_numpy(d).to(DEV)
    aabb = torch.tensor([-bound] * 3 + [bound] * 3, dtype=torch.float32, device=DEV)
    g = torch.from_numpy(grid).to(DEV)
    # utilities (raymarching.py:19-126)
    na, fa = ref.near_far_from_aabb(ro, rd, aabb, 0.2)
    nb, fb = mine.near_far_from_aabb(ro, rd, aabb, 0.2)
    assert torch.equal(na, nb) and torch.equal(fa, fb)
    coords = torch.randint(0, H, (5000, 3), device=DEV, dtype=torch.int32)
    ia, ib = ref.morton3D(coords), mine.morton3D(coords)
    assert torch.equal(ia, ib) and torch.equal(ref.morton3D_invert(ia), coords) and torch.equal(mine.morton3D_invert(ib), coords)
    bits_a, bits_b = ref.packbits(g, 0.01), mine.packbits(g, 0.01)
    assert torch.equal(bits_a, bits_b)
    assert torch.equal(ref.sph_from_ray(ro, rd, 8.0), mine.sph_from_ray(ro, rd, 8.0))
    # training march + composite (raymarching.py:161-297), both bookkeeping modes of the reference's wrapper
    for force_all, mean_count in ((True, -1), (False, 4096 * 64)):
        ca = torch.zeros(16, 2, dtype=torch.int32, device=DEV)
        cb = torch.zeros(16, 2, dtype=torch.int32, device=DEV)
        xa, da, la, ra = ref.march_rays_train(ro, rd, bound, bits_a, C, H, na, fa, ca[0], mean_count, False, 128, force_all, 1 / 128, 512)
        xb, db, lb, rb = mine.march_rays_train(ro, rd, bound, bits_b, C, H, nb, fb, cb[0], mean_count, False, 128, force_all, 1 / 128, 512)
        assert torch.equal(ca, cb) and xa.shape == xb.shape
        assert torch.equal(ra, rb) and torch.equal(xa, xb) and torch.equal(da, db) and torch.equal(la, lb)
        M = xa.shape[0]
        sig = torch.from_numpy(scenes.density(xa.cpu().numpy())).to(DEV).requires_grad_(True)
        rgb = torch.from_numpy(scenes.colour(xa.cpu().numpy())).to(DEV).requires_grad_(True)
        sig2, rgb2 = sig.detach().clone().requires_grad_(True), rgb.detach().clone().requires_grad_(True)
        wa, depa, ima = ref.composite_rays_train(sig, rgb, la, ra, 1e-4)
        wb, depb, imb = mine.composite_rays_train(sig2, rgb2, lb, rb, 1e-4)
        assert torch.equal(wa, wb) and torch.equal(depa, depb) and torch.equal(ima, imb) and M > 0
        gi = torch.randn_like(ima)
        (ima * gi).sum().backward(); (imb * gi).sum().backward()
        assert torch.equal(sig.grad, sig2.grad) and torch.equal(rgb.grad, rgb2.grad)
    # inference loop (raymarching.py:299-373), lock-step
    N = ro.shape[0]
    st = []
    for mod in (ref, mine):
        alive = torch.arange(N, dtype=torch.int32, device=DEV)
        rays_t = na.clone()
        ws, depth, image = torch.zeros(N, device=DEV), torch.zeros(N, device=DEV), torch.zeros(N, 3, device=DEV)
        n_alive, step = N, 0
        while step < 64 and n_alive > 0:
            n_step = max(min(N // n_alive, 8), 1)
            xyzs, dirs, deltas = mod.march_rays(n_alive, n_step, alive, rays_t, ro, rd, bound, bits_a, C, H, na, fa, 128, False, 1 / 128, 512)
            s_ = torch.from_numpy(scenes.density(xyzs.cpu().numpy())).to(DEV)
            c_ = torch.from_numpy(scenes.colour(xyzs.cpu().numpy())).to(DEV)
            mod.composite_rays(n_alive, n_step, alive, rays_t, s_, c_, deltas, ws, depth, image, 1e-2)
            alive = alive[alive >= 0]
            n_alive = alive.shape[0]
            step += n_step
        st.append((alive.clone(), rays_t.clone(), ws, depth, image))
    for u, v in zip(*st):
        assert torch.equal(u, v)
    assert float(st[0][2].max()) > 0.5
