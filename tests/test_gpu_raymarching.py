"""GPU parity tests of the ray-marching ops (raymarching.* -> C ABI -> sm_100a kernels) against
the CPU oracle.  Indices, counts, sample positions, deltas and bitfields: BIT-EXACT.  Composited
colour / depth / gradients: 1e-5 relative (fp32; the kernels use __expf like the reference)."""
import numpy as np
import pytest
import torch

import scenes

pytestmark = pytest.mark.gpu
DEV = "cuda"


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.fixture(scope="module")
def scene(oracle):
    out = {}
    for name, (H, C, bound, n) in {"c1": (128, 1, 1.0, 4096), "c2": (64, 2, 2.0, 2048)}.items():
        grid = scenes.density_grid(H=H, cascade=C, bound=bound)
        bits = oracle.packbits(grid, 0.01)
        o, d = scenes.camera_rays(n, seed=5, radius=1.6 * bound)
        aabb = np.array([-bound] * 3 + [bound] * 3, np.float32)
        nears, fars = oracle.near_far_from_aabb(o, d, aabb, 0.2)
        out[name] = dict(grid=grid, bits=bits, o=o, d=d, aabb=aabb, nears=nears, fars=fars, H=H, C=C, bound=bound)
    return out


def test_utils_bit_exact(oracle, scene):
    import raymarching as rm
    s = scene["c1"]
    n, f = rm.near_far_from_aabb(T(s["o"]), T(s["d"]), T(s["aabb"]), 0.2)
    assert np.array_equal(n.cpu().numpy(), s["nears"]) and np.array_equal(f.cpu().numpy(), s["fars"])
    assert (s["nears"] == np.finfo(np.float32).max).any() or True
    rng = np.random.default_rng(0)
    c = rng.integers(0, 1024, size=(100001, 3)).astype(np.int32)
    m = rm.morton3D(T(c))
    assert np.array_equal(m.cpu().numpy(), oracle.morton3D(c))
    assert np.array_equal(rm.morton3D_invert(m).cpu().numpy(), c)
    co = rm.sph_from_ray(T(s["o"]), T(s["d"]), 4.0).cpu().numpy()
    assert np.allclose(co, oracle.sph_from_ray(s["o"], s["d"], 4.0), atol=1e-5)


@pytest.mark.parametrize("shape", [(1, 128 ** 3), (2, 64 ** 3), (1, 8 * 13), (3, 8 * 5)])
def test_packbits_bit_exact(oracle, shape):
    import raymarching as rm
    rng = np.random.default_rng(0)
    g = rng.standard_normal(shape).astype(np.float32)
    g.reshape(-1)[:40] = 0.3
    g.reshape(-1)[40:48] = -1.0  # untrained cells
    bits = rm.packbits(T(g), 0.3)
    assert bits.dtype == torch.uint8 and bits.numel() == g.size // 8
    assert np.array_equal(bits.cpu().numpy(), oracle.packbits(g, 0.3))
    assert np.array_equal(bits.cpu().numpy(), np.packbits(g.reshape(-1) > 0.3, bitorder="little"))
    # in-place variant on an unaligned view exercises the byte-granular tail kernel
    buf = torch.zeros(g.size // 8 + 1, dtype=torch.uint8, device=DEV)
    rm.packbits(T(g), 0.3, buf[1:])
    assert np.array_equal(buf[1:].cpu().numpy(), oracle.packbits(g, 0.3))


@pytest.fixture(params=[(1, 1), (1, 0), (0, 0)], ids=["warp_per_ray_masks", "warp_per_ray_remarch", "thread_per_ray"])
def variant(request):
    """Every implementation of the training marcher / compositor must satisfy the same contract."""
    import ngp_abi as abi
    abi.set_option("march_variant", request.param[0])
    abi.set_option("composite_variant", request.param[0])
    abi.set_option("march_masks", request.param[1])
    yield request.param
    abi.set_option("march_variant", -1)
    abi.set_option("composite_variant", -1)
    abi.set_option("march_masks", 1)


@pytest.mark.parametrize("name,dt_gamma,perturb", [("c1", 0.0, False), ("c1", 1.0 / 128, True), ("c2", 0.0, True), ("c2", 1.0 / 128, False)])
def test_march_rays_train_bit_exact(oracle, scene, name, dt_gamma, perturb, variant):
    import raymarching as rm
    s = scene[name]
    N = s["o"].shape[0]
    M = N * 1024
    noises = np.random.default_rng(1).random(N).astype(np.float32) if perturb else np.zeros(N, np.float32)
    ref = oracle.march_rays_train(s["o"], s["d"], s["bound"], s["bits"], s["C"], s["H"], s["nears"], s["fars"], M, noises, dt_gamma, 1024)
    be = rm._backend
    xyzs, dirs, deltas = torch.zeros(M, 3, device=DEV), torch.zeros(M, 3, device=DEV), torch.zeros(M, 2, device=DEV)
    rays = torch.empty(N, 3, dtype=torch.int32, device=DEV)
    counter = torch.zeros(2, dtype=torch.int32, device=DEV)
    be.march_rays_train(T(s["o"]), T(s["d"]), T(s["bits"]), s["bound"], dt_gamma, 1024, N, s["C"], s["H"], M, T(s["nears"]), T(s["fars"]),
                        xyzs, dirs, deltas, rays, counter, T(noises))
    assert np.array_equal(counter.cpu().numpy(), ref[4]) and ref[4][0] > N
    assert np.array_equal(rays.cpu().numpy(), ref[3])
    assert np.array_equal(xyzs.cpu().numpy(), ref[0])
    assert np.array_equal(dirs.cpu().numpy(), ref[1])
    assert np.array_equal(deltas.cpu().numpy(), ref[2])
    # a second call accumulates into the counter like the reference's atomicAdd (:405-406)
    be.march_rays_train(T(s["o"]), T(s["d"]), T(s["bits"]), s["bound"], dt_gamma, 1024, N, s["C"], s["H"], M, T(s["nears"]), T(s["fars"]),
                        xyzs, dirs, deltas, rays, counter, T(noises))
    assert counter[1].item() == 2 * N and counter[0].item() == 2 * ref[4][0]


def test_march_rays_train_budget_and_api(oracle, scene, variant):
    import raymarching as rm
    s = scene["c1"]
    N = s["o"].shape[0]
    o, d, bits, nears, fars = T(s["o"]), T(s["d"]), T(s["bits"]), T(s["nears"]), T(s["fars"])
    # first-steps mode: full buffer, count read back, trimmed past the next multiple of `align`
    counter = torch.zeros(2, dtype=torch.int32, device=DEV)
    xyzs, dirs, deltas, rays = rm.march_rays_train(o, d, s["bound"], bits, s["C"], s["H"], nears, fars, counter, -1, False, 128, False, 0, 1024)
    total = counter[0].item()
    assert xyzs.shape[0] == total + 128 - total % 128 and xyzs.shape == dirs.shape and deltas.shape[1] == 2
    # steady-state mode with an under-estimated budget: overflowing rays are dropped, the rest are bit-identical
    mean_count = total // 2
    counter.zero_()
    x2, d2, l2, r2 = rm.march_rays_train(o, d, s["bound"], bits, s["C"], s["H"], nears, fars, counter, mean_count, False, 128, False, 0, 1024)
    M = mean_count + 128 - mean_count % 128
    assert x2.shape[0] == M and counter[0].item() == total
    ref = oracle.march_rays_train(s["o"], s["d"], s["bound"], s["bits"], s["C"], s["H"], s["nears"], s["fars"], M)
    assert np.array_equal(x2.cpu().numpy(), ref[0]) and np.array_equal(l2.cpu().numpy(), ref[2]) and np.array_equal(r2.cpu().numpy(), ref[3])
    # empty input
    z = torch.zeros(0, 3, device=DEV)
    out = rm.march_rays_train(z, z, 1.0, bits, 1, 128, torch.zeros(0, device=DEV), torch.zeros(0, device=DEV), None, 64, False, 128)
    assert out[3].shape == (0, 3)


def test_composite_train_forward_backward(oracle, scene, variant):
    import raymarching as rm
    s = scene["c1"]
    N = 2048
    ref = oracle.march_rays_train(s["o"][:N], s["d"][:N], s["bound"], s["bits"], s["C"], s["H"], s["nears"][:N], s["fars"][:N], N * 1024)
    m = int(ref[4][0])
    m_pad = m + 128 - m % 128
    xyzs, deltas, rays = ref[0][:m_pad], ref[2][:m_pad], ref[3]
    rng = np.random.default_rng(0)
    sig = (scenes.density(xyzs) * rng.random(m_pad)).astype(np.float32)   # mixes early termination and long segments
    rgb = scenes.colour(xyzs)
    for T_thresh in (1e-4, 0.0):
        ws_r, dep_r, img_r = oracle.composite_rays_train_forward(sig, rgb, deltas, rays, T_thresh)
        sg, cl = T(sig).requires_grad_(True), T(rgb).requires_grad_(True)
        ws, dep, img = rm.composite_rays_train(sg, cl, T(deltas), T(rays), T_thresh)
        assert np.allclose(ws.detach().cpu().numpy(), ws_r, rtol=1e-5, atol=1e-6)
        assert np.allclose(dep.detach().cpu().numpy(), dep_r, rtol=1e-5, atol=1e-5)
        assert np.allclose(img.detach().cpu().numpy(), img_r, rtol=1e-5, atol=1e-6)
        g_ws, g_dep, g_img = rng.standard_normal(N).astype(np.float32), rng.standard_normal(N).astype(np.float32), rng.standard_normal((N, 3)).astype(np.float32)
        torch.autograd.backward([ws, dep, img], [T(g_ws), T(g_dep), T(g_img)])
        gs_r, gc_r = oracle.composite_rays_train_backward(g_ws, g_dep, g_img, sig, rgb, deltas, rays, ws_r, dep_r, img_r, T_thresh)
        assert np.allclose(sg.grad.cpu().numpy(), gs_r, rtol=1e-4, atol=1e-5 * np.abs(gs_r).max())
        assert np.allclose(cl.grad.cpu().numpy(), gc_r, rtol=1e-5, atol=1e-6)


def test_inference_loop_bit_exact_and_compaction(oracle, scene):
    """The reference's eval loop (nerf/renderer.py:394-420) run twice: host compaction via the oracle and
    device compaction via compact_alive; samples bit-exact per iteration, images within tolerance."""
    import raymarching as rm
    s = scene["c1"]
    N = s["o"].shape[0]
    o, d, bits, nears, fars = T(s["o"]), T(s["d"]), T(s["bits"]), T(s["nears"]), T(s["fars"])
    ws, dep, img = torch.zeros(N, device=DEV), torch.zeros(N, device=DEV), torch.zeros(N, 3, device=DEV)
    ws_r, dep_r, img_r = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, 3), np.float32)
    alive = torch.arange(N, dtype=torch.int32, device=DEV)
    alive_r = np.arange(N, dtype=np.int32)
    rays_t, rays_t_r = nears.clone(), s["nears"].copy()
    step, iters = 0, 0
    while step < 1024:
        n_alive = alive_r.size
        if n_alive == 0:
            break
        n_step = max(min(N // n_alive, 8), 1)
        xr, dr, lr = oracle.march_rays(n_alive, n_step, alive_r, rays_t_r, s["o"], s["d"], s["bound"], s["bits"], s["C"], s["H"], s["nears"], s["fars"])
        x, dd, dl = rm.march_rays(n_alive, n_step, alive, rays_t, o, d, s["bound"], bits, s["C"], s["H"], nears, fars, 128, False, 0, 1024)
        M = n_alive * n_step
        assert x.shape[0] == M + 128 - M % 128
        assert np.array_equal(x[:M].cpu().numpy(), xr) and np.array_equal(dl[:M].cpu().numpy(), lr) and np.array_equal(dd[:M].cpu().numpy(), dr)
        sig, rgb = scenes.density(xr), scenes.colour(xr)
        oracle.composite_rays(n_alive, n_step, alive_r, rays_t_r, sig, rgb, lr, ws_r, dep_r, img_r, 1e-4)
        sg, cl = torch.zeros(x.shape[0], device=DEV), torch.zeros(x.shape[0], 3, device=DEV)
        sg[:M], cl[:M] = T(sig), T(rgb)
        rm.composite_rays(n_alive, n_step, alive, rays_t, sg, cl, dl, ws, dep, img, 1e-4)
        assert np.array_equal(alive[:n_alive].cpu().numpy(), alive_r)          # same rays terminated
        assert np.allclose(rays_t.cpu().numpy(), rays_t_r, rtol=1e-6, atol=1e-6)
        # keep the two loops in lock-step: resume both from the oracle's t
        rays_t = T(rays_t_r)
        comp, n_out = rm.compact_alive(alive, n_alive)
        alive_r = np.ascontiguousarray(alive_r[alive_r >= 0])
        assert n_out.item() == alive_r.size and np.array_equal(comp[:alive_r.size].cpu().numpy(), alive_r)
        alive = comp
        step += n_step
        iters += 1
    assert iters > 4
    assert np.allclose(img.cpu().numpy(), img_r, rtol=1e-4, atol=1e-5) and np.allclose(ws.cpu().numpy(), ws_r, rtol=1e-4, atol=1e-5)
    assert np.allclose(dep.cpu().numpy(), dep_r, rtol=1e-4, atol=1e-4)


def test_compact_alive_large():
    import raymarching as rm
    torch.manual_seed(0)
    n = 1920 * 1080
    a = torch.arange(n, dtype=torch.int32, device=DEV)
    a[torch.rand(n, device=DEV) < 0.37] = -1
    out, cnt = rm.compact_alive(a)
    ref = a[a >= 0]
    assert cnt.item() == ref.numel() and torch.equal(out[: ref.numel()], ref)
    out2, cnt2 = rm.compact_alive(torch.full((1000,), -1, dtype=torch.int32, device=DEV))
    assert cnt2.item() == 0


def test_raymarching_error_behaviour(scene):
    import raymarching as rm
    be = rm._backend
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        be.packbits(torch.zeros(64), 8, 0.0, torch.zeros(8, dtype=torch.uint8, device=DEV))
    with pytest.raises(RuntimeError, match="must be a uint8 tensor"):
        be.packbits(torch.zeros(64, device=DEV), 8, 0.0, torch.zeros(8, dtype=torch.int32, device=DEV))
    with pytest.raises(RuntimeError, match="expected at least"):
        be.packbits(torch.zeros(32, device=DEV), 8, 0.0, torch.zeros(8, dtype=torch.uint8, device=DEV))
    # CPU inputs to the public ops are moved to the GPU like the reference does (raymarching.py:34-35)
    s = scene["c1"]
    n, f = rm.near_far_from_aabb(torch.from_numpy(s["o"][:10]), torch.from_numpy(s["d"][:10]), T(s["aabb"]), 0.2)
    assert n.is_cuda and np.array_equal(n.cpu().numpy(), s["nears"][:10])


@pytest.mark.parametrize("max_steps,dt_gamma", [(40, 0.0), (100, 1.0 / 64), (1024, 0.0)])
def test_march_rays_train_step_cap_and_full_occupancy(oracle, scene, variant, max_steps, dt_gamma):
    """Fully occupied grid: every ray hits the max_steps cap (the `step < max_steps` half of the loop condition),
    runs of occupied points span whole 32-point chunks, and nothing is skipped."""
    import raymarching as rm
    s = scene["c2"]
    N = 1000
    bits = np.full_like(s["bits"], 0xFF)
    bits[::7] = 0x5A  # a few holes so that skips and runs interleave
    noises = np.random.default_rng(5).random(N).astype(np.float32)
    M = N * max_steps
    ref = oracle.march_rays_train(s["o"][:N], s["d"][:N], s["bound"], bits, s["C"], s["H"], s["nears"][:N], s["fars"][:N], M, noises, dt_gamma, max_steps)
    xyzs, dirs, deltas = torch.zeros(M, 3, device=DEV), torch.zeros(M, 3, device=DEV), torch.zeros(M, 2, device=DEV)
    rays = torch.empty(N, 3, dtype=torch.int32, device=DEV)
    counter = torch.zeros(2, dtype=torch.int32, device=DEV)
    rm._backend.march_rays_train(T(s["o"][:N]), T(s["d"][:N]), T(bits), s["bound"], dt_gamma, max_steps, N, s["C"], s["H"], M, T(s["nears"][:N]),
                                 T(s["fars"][:N]), xyzs, dirs, deltas, rays, counter, T(noises))
    assert np.array_equal(counter.cpu().numpy(), ref[4])
    assert (ref[3][:, 2] == max_steps).any()
    assert np.array_equal(rays.cpu().numpy(), ref[3])
    assert np.array_equal(xyzs.cpu().numpy(), ref[0])
    assert np.array_equal(deltas.cpu().numpy(), ref[2])


def test_mark_untrained_grid_matches_oracle_and_stays_unoccupied():
    """ngp_mark_untrained_grid (one kernel) against the numpy restatement of nerf/renderer.py:433-496: per-cell camera
    counts bit-exact, unseen cells -1; a density-grid update afterwards never turns them on (renderer.py:565-566)."""
    import raymarching
    from oracle import cpu_ref
    from test_oracle import _orbit_poses
    cpu_ref.build()
    H, C, bound = 32, 2, 2.0
    poses = _orbit_poses(70)  # > 64: two shared-memory batches of cameras
    intr = (55.0, 50.0, 20.0, 16.0)
    grid = torch.rand(C, H ** 3, device="cuda")
    before = grid.clone()
    count = raymarching.mark_untrained_grid(poses, intr, grid, bound, return_count=True)
    ref = cpu_ref.mark_untrained_grid(poses, intr, bound, C, H)
    assert np.array_equal(count.cpu().numpy(), ref)
    unseen = torch.from_numpy(ref == 0).cuda()
    assert 0 < int(unseen.sum()) < unseen.numel()
    assert torch.equal(grid[unseen], torch.full_like(grid[unseen], -1)) and torch.equal(grid[~unseen], before[~unseen])
    # the trainer's update keeps them out of the bitfield
    import nerf
    tr = nerf.NGPTrainer(bound=1.0, num_rays=256, use_graph=False, log2_hashmap_size=15)
    n_marked = tr.mark_untrained_grid(_orbit_poses(6), (300.0, 300.0, 64.0, 64.0))
    assert 0 < n_marked < tr.density_grid.numel()
    marked = tr.density_grid < 0
    tr.update_extra_state()
    assert torch.equal(tr.density_grid < 0, marked)
    bits = tr.density_bitfield.cpu().numpy()
    occ = np.unpackbits(bits, bitorder="little").astype(bool)
    assert not occ[marked.view(-1).cpu().numpy()].any()
