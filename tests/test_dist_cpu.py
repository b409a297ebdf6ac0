"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: the data-parallel gradient exchange of the training
step and the row-band sharding of inference frames.  The per-rank gradients come from the CPU oracle of the step."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "torch-ngp_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _problem():
    from oracle import cpu_ref, nerf_step_ref as R
    import scenes
    cpu_ref.build()
    offs, pls = cpu_ref.make_offsets(num_levels=16, log2_hashmap_size=10, desired_resolution=256)
    rng = np.random.default_rng(0)
    k = np.sqrt(3 / 64)
    st = R.RefState(((rng.random((offs[-1], 2)) * 2 - 1) * 1e-1).astype(np.float32), ((rng.random(64 * 112) * 2 - 1) * k).astype(np.float32),
                    ((rng.random(64 * 176) * 2 - 1) * k).astype(np.float32), offs, pls, 16)
    H = 32
    bits = scenes.bitfield_from_grid(scenes.density_grid(H=H), 0.01)
    ro, rd = scenes.camera_rays(64, seed=3)
    gt = rng.random((64, 3)).astype(np.float32)
    return R, st, bits, H, ro, rd, gt


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nerf.parallel import allreduce_gradients, row_band
        R, st, bits, H, ro, rd, gt = _problem()
        n = ro.shape[0] // world
        sl = slice(rank * n, (rank + 1) * n)
        loss, grads, _ = R.forward_backward(st, ro[sl], rd[sl], gt[sl], bits, 1, H, n * 256, max_steps=256)
        g16 = torch.from_numpy(grads[0]).half()
        g32 = torch.from_numpy(np.concatenate([grads[1], grads[2]]))
        allreduce_gradients(g16, g32)
        R.adam_step(st, (g16.float().numpy(), g32[:64 * 112].numpy(), g32[64 * 112:].numpy()), grad_div=world)
        bands = [None] * world
        dist.all_gather_object(bands, row_band(1080, rank, world))
        if rank == 0:
            torch.save({"table": st.table, "sigma_w": st.sigma_w, "color_w": st.color_w, "bands": bands, "g32": g32.numpy()}, out)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_data_parallel_step_equals_global_batch_step(tmp_path):
    world, port, out = 2, 29500 + os.getpid() % 2000, str(tmp_path / "dp.pt")
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    got = torch.load(out, weights_only=False)
    # single process, the whole batch at once
    R, st, bits, H, ro, rd, gt = _problem()
    n = ro.shape[0] // world
    # per-rank sample budgets (n * 256) never overflow here, so marching the halves equals marching the whole
    loss, grads, _ = R.forward_backward(st, ro, rd, gt, bits, 1, H, ro.shape[0] * 256, max_steps=256)
    g32 = np.concatenate([grads[1], grads[2]])
    # sum over ranks of (mean over n rays) = world * (mean over the global batch)
    assert np.allclose(got["g32"], world * g32, rtol=2e-3, atol=2e-3 * np.abs(g32).max())
    R.adam_step(st, grads)
    assert np.allclose(got["sigma_w"], st.sigma_w, rtol=0, atol=2e-3)  # lr 1e-2 * sign-like updates; fp16 table gradient rounding
    assert np.abs(got["table"] - st.table).mean() < 1e-3
    assert got["bands"] == [(0, 540), (540, 1080)]


def test_row_stripes_cover_the_frame():
    import torch
    from nerf.parallel import row_stripes
    for h, w, st in [(1080, 8, 8), (1080, 7, 16), (5, 8, 8), (720, 1, 8), (1081, 4, 3)]:
        parts = [row_stripes(h, r, w, st) for r in range(w)]
        allr = torch.cat(parts)
        assert allr.numel() == h and torch.equal(torch.sort(allr).values, torch.arange(h))
        assert max(p.numel() for p in parts) - min(p.numel() for p in parts) <= st


def test_row_bands_cover_the_frame():
    from nerf.parallel import row_band, rays_per_rank
    for h, w in ((1080, 8), (1080, 7), (5, 8), (800, 1)):
        bands = [row_band(h, r, w) for r in range(w)]
        assert bands[0][0] == 0 and bands[-1][1] == h
        assert all(a[1] == b[0] for a, b in zip(bands, bands[1:]))
        sizes = [b - a for a, b in bands]
        assert max(sizes) - min(sizes) <= 1
    assert rays_per_rank(1 << 18, 8) == 32768
