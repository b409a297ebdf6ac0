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


def _transcribed_update(grid, tmp, decay, density_thresh, packbits):
    """nerf/renderer.py:582-590 with torch ops, as the reference executes them; packbits = the op to end with."""
    grid = grid.clone()
    valid_mask = (grid >= 0) & (tmp >= 0)
    grid[valid_mask] = torch.maximum(grid[valid_mask] * decay, tmp[valid_mask])
    mean_density = torch.mean(grid.clamp(min=0)).item()
    return grid, mean_density, packbits(grid, min(mean_density, density_thresh))


@pytest.mark.parametrize("cascade,H,density_thresh", [(1, 128, 10.0), (3, 64, 10.0), (2, 64, 0.2)])
def test_density_grid_update_kernels_vs_transcription(oracle, cascade, H, density_thresh):
    """R10: ngp_density_ema + ngp_packbits_mean (mean and threshold stay on the device) against (a) the torch transcription of
    renderer.py:582-590 ending in the REFERENCE extension's packbits when it is built (else this library's, itself pinned
    bit-exact to it), (b) the oracle twin.  Grid bit-exact.  Mean: the kernel carries the sum in double, so its fp32 mean is the
    correctly rounded one (== the oracle's float64 evaluation); torch.mean's fp32 tree sum may sit a few ulp away (2e-7).
    Bitfield bit-exact on every cell that is not inside that ulp-wide window around the threshold (at most a handful of 2 M)."""
    import ngp_abi as abi
    import raymarching
    from oracle import ref_ext
    ref = ref_ext.load("_raymarching")

    def pack(grid, thresh):
        bits = torch.empty(grid.numel() // 8, dtype=torch.uint8, device=DEV)
        if ref is not None:
            ref.packbits(grid.contiguous(), grid.numel() // 8, thresh, bits)
            return bits
        return raymarching.packbits(grid, thresh)

    torch.manual_seed(cascade * 1000 + H)
    n = H ** 3
    grid = torch.rand(cascade, n, device=DEV) ** 4 * 30
    grid[torch.rand(cascade, n, device=DEV) < 0.2] = -1.0   # untrained cells (mark_untrained_grid)
    grid[torch.rand(cascade, n, device=DEV) < 0.1] = 0.0
    tmp = torch.rand(cascade, n, device=DEV) ** 4 * 30
    tmp[torch.rand(cascade, n, device=DEV) < 0.5] = -1.0    # cells the partial update did not probe
    tmp[torch.rand(cascade, n, device=DEV) < 0.05] = 0.0
    for it in range(3):  # the EMA is stateful: a few rounds
        g_ref, mean_ref, bits_ref = _transcribed_update(grid, tmp, 0.95, density_thresh, pack)
        g_orc, mean_orc, bits_orc = oracle.density_grid_update(grid.cpu().numpy(), tmp.cpu().numpy(), 0.95, density_thresh)
        g = grid.clone()
        msum = torch.zeros(1, dtype=torch.float64, device=DEV)
        bits = torch.zeros(cascade * n // 8, dtype=torch.uint8, device=DEV)
        abi.check(abi.lib.ngp_density_ema(g.data_ptr(), tmp.data_ptr(), 0.95, g.numel(), msum.data_ptr(), abi.stream()))
        abi.check(abi.lib.ngp_packbits_mean(g.data_ptr(), g.numel() // 8, msum.data_ptr(), 1.0 / g.numel(), density_thresh, bits.data_ptr(), abi.stream()))
        mean = float(np.float32(float(msum.item()) / g.numel()))
        assert torch.equal(g, g_ref) and np.array_equal(g.cpu().numpy(), g_orc)
        assert mean == mean_orc                          # correctly rounded fp32 mean, whatever the addition order
        assert abs(mean - mean_ref) <= 3e-7 * mean_ref   # torch.mean's own fp32 reduction
        assert np.array_equal(bits.cpu().numpy(), bits_orc)  # same mean -> same threshold -> bit-exact
        thr = min(mean_ref, density_thresh)
        far = ((g - thr).abs() > 1e-6 * thr).view(-1).cpu().numpy()   # cells torch.mean's last bits cannot flip
        assert (~far).sum() <= 8
        occ = np.unpackbits(bits.cpu().numpy(), bitorder="little").astype(bool)
        assert np.array_equal(occ[far], np.unpackbits(bits_ref.cpu().numpy(), bitorder="little").astype(bool)[far])
        assert not occ[(g < 0).view(-1).cpu().numpy()].any() and 0 < occ.sum() < occ.size
        grid = g
        tmp = torch.where(torch.rand_like(tmp) < 0.5, torch.rand_like(tmp) * 20, torch.full_like(tmp, -1.0))


@pytest.mark.parametrize("bound", [1.0, 2.0])
def test_update_extra_state_vs_transcription(bound):
    """NGPTrainer.update_extra_state against a transcription of nerf/renderer.py:498-598 that draws the same torch random
    numbers in the same order and queries the same density function, for the full sweep (iter_density < 16) and the partial
    update (>= 16), 1 and 2 cascades.  Full sweep: every cell is written once -> density grid bit-exact.  Partial update: the
    reference draws cells WITH duplicates (:554 "allow for duplication") and `tmp_grid[cas, indices] = sigmas` lets an
    arbitrary one of them win, in the reference as here; cells probed once are bit-exact, cells probed several times must
    hold one of their candidates' results (checked as min <= value <= max over the candidates).  The bitfield is the
    reference packbits of the resulting grid at min(mean, density_thresh)."""
    import nerf
    import raymarching
    import ngp_abi as abi
    H = 32
    tr = nerf.NGPTrainer(bound=bound, num_rays=256, use_graph=False, log2_hashmap_size=14, grid_size=H, density_thresh=0.9)
    with torch.no_grad():  # a table with structure, so that densities spread around the threshold
        tr.embeddings.copy_(torch.randn_like(tr.embeddings) * 0.5)
        abi.check(abi.lib.ngp_cast_f16(tr.params.data_ptr(), tr.table16.data_ptr(), tr.n16, abi.stream()))

    def reference_probes(density_grid, iter_density):
        """renderer.py:506-574 -> per cascade (indices, sigmas) in the order the reference assigns them"""
        probes = []
        if iter_density < 16:
            X = torch.arange(H, dtype=torch.int32, device=DEV).split(128)
            for xs in X:
                for ys in X:
                    for zs in X:
                        xx, yy, zz = torch.meshgrid(xs, ys, zs, indexing="ij")
                        coords = torch.cat([xx.reshape(-1, 1), yy.reshape(-1, 1), zz.reshape(-1, 1)], dim=-1)
                        indices = raymarching.morton3D(coords).long()
                        xyzs = 2 * coords.float() / (H - 1) - 1
                        for cas in range(tr.cascade):
                            b = min(2 ** cas, tr.bound)
                            half_grid_size = b / H
                            cas_xyzs = xyzs * (b - half_grid_size)
                            cas_xyzs += (torch.rand_like(cas_xyzs) * 2 - 1) * half_grid_size
                            probes.append((cas, indices, tr.density(cas_xyzs)))
        else:
            N = H ** 3 // 4
            for cas in range(tr.cascade):
                coords = torch.randint(0, H, (N, 3), device=DEV)
                indices = raymarching.morton3D(coords).long()
                occ_indices = torch.nonzero(density_grid[cas] > 0).squeeze(-1)
                rand_mask = torch.randint(0, occ_indices.shape[0], [N], dtype=torch.long, device=DEV)
                occ_indices = occ_indices[rand_mask]
                occ_coords = raymarching.morton3D_invert(occ_indices)
                indices = torch.cat([indices, occ_indices], dim=0)
                coords = torch.cat([coords, occ_coords], dim=0)
                xyzs = 2 * coords.float() / (H - 1) - 1
                b = min(2 ** cas, tr.bound)
                half_grid_size = b / H
                cas_xyzs = xyzs * (b - half_grid_size)
                cas_xyzs += (torch.rand_like(cas_xyzs) * 2 - 1) * half_grid_size
                probes.append((cas, indices, tr.density(cas_xyzs)))
        return probes

    for iter_density in (0, 1, 16, 17):
        tr.iter_density = iter_density
        before = tr.density_grid.clone()
        torch.manual_seed(100 + iter_density)
        probes = reference_probes(before, iter_density)
        torch.manual_seed(100 + iter_density)
        tr.update_extra_state()
        assert tr.iter_density == iter_density + 1
        # renderer.py:582-583 per candidate: a probed cell ends as max(old * decay, sigma) (old >= 0, sigma >= 0), else unchanged
        assert len(probes) == tr.cascade  # H = 32 < S = 128: one block, one probe batch per cascade
        lo, hi, cnt = before.clone(), before.clone(), torch.zeros_like(before)
        for cas, idx, sig in probes:
            old = before[cas, idx]
            res = torch.where((old >= 0) & (sig >= 0), torch.maximum(old * 0.95, sig), old)
            cnt[cas].index_add_(0, idx, torch.ones_like(sig))
            hi[cas].scatter_reduce_(0, idx, res, "amax", include_self=False)  # untouched cells keep `before`
            lo[cas].scatter_reduce_(0, idx, res, "amin", include_self=False)
        g = tr.density_grid
        once = cnt <= 1
        assert torch.equal(g[once], lo[once]) and torch.equal(lo[once], hi[once]), iter_density      # bit-exact where defined
        assert bool(((g >= lo) & (g <= hi)).all()), iter_density                                    # one of the candidates elsewhere
        if iter_density < 16:
            assert bool(once.all())
        else:
            assert 0.02 < float((~once).float().mean()) < 0.9
        # renderer.py:584-590 on the grid that resulted
        mean = float(np.float32(g.double().clamp(min=0).mean().item()))
        assert tr.mean_density() == mean
        bits_ref = raymarching.packbits(g, min(mean, tr.density_thresh))
        assert torch.equal(tr.density_bitfield, bits_ref)
        occ = np.unpackbits(bits_ref.cpu().numpy(), bitorder="little")
        assert 0 < occ.sum() < occ.size
