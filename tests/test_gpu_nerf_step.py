"""Fused training step (nerf/trainer.py on raw buffers through the C ABI) against the autograd path through the
plugin modules (NeRFNetwork = GridEncoder + FFMLP + SHEncoder, raymarching.composite_rays_train, torch Adam +
GradScaler semantics), i.e. against what /root/reference/nerf/utils.py::train_step + nerf/renderer.py::run_cuda compute.

Tolerances: the sample buffers (march) are bit-exact; activations are fp16, table gradients are accumulated with
fp16 atomics in a different order, so gradients are compared in relative Frobenius norm (1e-2) and the loss to 1e-4."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make(num_rays=2048, graph=False, **kw):
    import nerf
    from nerf import synthetic
    tr = nerf.NGPTrainer(bound=1.0, num_rays=num_rays, use_graph=graph, log2_hashmap_size=17, **kw)
    o, d, gt = synthetic.make_pool(4, num_rays, "cuda", seed=3)
    tr.set_pool(o, d, gt)
    return tr


def _twin(tr):
    import nerf
    net = nerf.NeRFNetwork(bound=tr.bound, log2_hashmap_size=17).cuda()
    with torch.no_grad():
        net.encoder.embeddings.copy_(tr.embeddings)
        net.sigma_net.weights.copy_(tr.sigma_w)
        net.color_net.weights.copy_(tr.color_w)
    net.train()
    return net


def _rel(a, b):
    return float((a.double() - b.double()).norm() / max(1e-30, b.double().norm()))


def test_forward_backward_matches_autograd_path():
    import raymarching
    tr = _make()
    tr.update_extra_state()
    assert tr.density_bitfield.any(), "density grid update left the occupancy empty"
    tr.next_pool_batch()
    net = _twin(tr)
    tr._forward_backward()
    torch.cuda.synchronize()
    n_samples = int(tr.step_counter[0, 0])
    assert 0 < n_samples <= tr.M, (n_samples, tr.M)
    scale = float(tr.state[4])

    # same samples through the autograd modules
    xyzs, dirs, deltas, rays = tr.xyzs.clone(), tr.dirs.clone(), tr.deltas.clone(), tr.rays.clone()
    with torch.autocast("cuda", dtype=torch.half):
        sigmas, rgbs = net(xyzs, dirs)
        ws, depth, image = raymarching.composite_rays_train(sigmas, rgbs, deltas, rays, tr.T_thresh)  # custom_fwd casts to fp32
    pred = image + (1 - ws).unsqueeze(-1) * tr.bg_color
    loss = torch.nn.functional.mse_loss(pred, tr.gt, reduction="none").mean(-1).mean()
    (loss * scale).backward()
    torch.cuda.synchronize()

    assert torch.equal(sigmas.detach()[:n_samples], tr.sigmas[:n_samples]) or _rel(tr.sigmas[:n_samples], sigmas.detach()[:n_samples]) < 1e-3
    assert _rel(tr.rgbs[:n_samples], rgbs.detach().float()[:n_samples]) < 1e-3
    assert _rel(tr.image, image.detach()) < 1e-4
    assert abs(float(tr.state[7]) - float(loss.detach())) < 1e-4 * max(1.0, float(loss.detach()))
    g_tab = net.encoder.embeddings.grad
    assert g_tab.abs().max() > 0
    assert _rel(tr.g16.float(), g_tab) < 1e-2, _rel(tr.g16.float(), g_tab)
    assert _rel(tr.g_sig, net.sigma_net.weights.grad) < 1e-2, _rel(tr.g_sig, net.sigma_net.weights.grad)
    assert _rel(tr.g_col, net.color_net.weights.grad) < 1e-2, _rel(tr.g_col, net.color_net.weights.grad)


def test_fused_nets_equal_the_unfused_chain():
    """The glue between the two MLPs fused into the tcgen05 kernels (ngp_nerf_sigma_forward / color_forward / color_backward:
    trunc_exp, SH4 + geo-feature tile built in shared memory, sigmoid, and their backward, network_ff.py:51-74) against the same
    step through ngp_ffmlp_* + ngp_nerf_mid_* + ngp_nerf_rgb_*: every per-sample tensor bit for bit; weight / table gradients up
    to the order of their atomic sums.  Also the inference renderer, which shades through the same two entry points."""
    from nerf import synthetic
    out = []
    for fused, merged in ((True, True), (False, False), (True, False)):  # one launch for both nets / the unfused chain / one launch per net
        torch.manual_seed(7)
        tr = _make(fused_nets=fused)
        tr.merged_forward = merged
        assert tr.fused_nets == fused
        tr.update_extra_state()
        tr.next_pool_batch()
        torch.manual_seed(8)
        tr._forward_backward()
        torch.cuda.synchronize()
        out.append(tr)
    a, b, c = out
    n = int(a.step_counter[0, 0])
    for name in ("h", "sigmas", "rgbs", "g_h", "g_enc"):
        assert torch.equal(getattr(a, name)[:n], getattr(c, name)[:n]), name
    assert n == int(b.step_counter[0, 0]) and n > 1000
    assert a.cin.shape[0] == 1 and b.cin.shape[0] == b.M  # the fused path does not even allocate the matrices between the nets
    for name in ("h", "sigmas", "rgbs", "g_sigmas", "g_rgbs", "g_h", "g_enc"):
        assert torch.equal(getattr(a, name)[:n], getattr(b, name)[:n]), name
    assert torch.equal(a.image, b.image) and abs(float(a.state[7]) - float(b.state[7])) <= 1e-6 * abs(float(b.state[7]))
    assert _rel(a.g_col, b.g_col) < 1e-5 and _rel(a.g_sig, b.g_sig) < 1e-5
    assert _rel(a.g16.float(), b.g16.float()) < 2e-2
    o, d = synthetic.frame_rays(96, 54, "cuda", pose_index=5)
    b.load_state_dict(a.state_dict())
    ia, da = a.render(o, d, use_graph=False)
    ib, db = b.render(o, d, use_graph=False)
    assert torch.equal(ia, ib) and torch.equal(da, db)


def test_host_batches_fed_on_a_side_stream_equal_device_batches():
    """train_step(pinned host tensors): the batch is staged on a side stream (nerf/feed.py) and the loss collected asynchronously
    (loss_async); same samples and losses as with device-resident batches, including across the staging ring's reuse."""
    from nerf import synthetic
    o, d, gt = synthetic.make_pool(6, 1024, "cuda", seed=5)
    host = [t.cpu().pin_memory() for t in (o, d, gt)]
    runs = []
    for use_host in (False, True):
        torch.manual_seed(3)
        tr = _make(num_rays=1024, graph=True, perturb=False, update_interval=0)
        tr.update_extra_state()
        counts, losses, pend = [], [], None
        for k in range(6):
            sl = slice(k * 1024, (k + 1) * 1024)
            src = host if use_host else (o, d, gt)
            tr.train_step(src[0][sl], src[1][sl], src[2][sl])
            counts.append(tr.step_counter[0, :1].clone())
            if pend is not None:
                losses.append(pend())
            pend = tr.loss_async()
        losses.append(pend())
        torch.cuda.synchronize()
        runs.append(([int(c) for c in counts], losses))
    assert runs[0][0] == runs[1][0] and runs[0][0][0] > 1000
    assert all(abs(a - b) <= 2e-2 * abs(b) + 1e-7 for a, b in zip(runs[1][1], runs[0][1])), runs
    assert abs(runs[1][1][0] - runs[0][1][0]) <= 1e-6 * runs[0][1][0]  # first step: identical parameters, identical batch


def test_optimizer_matches_torch_adam_and_grad_scaler():
    tr = _make(num_rays=1024)
    tr.update_extra_state()
    p_ref = tr.params.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1e-2, betas=(0.9, 0.99), eps=1e-15)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda it: 0.1 ** min(it / 30000, 1))
    for step in range(3):
        tr.next_pool_batch()
        tr._forward_backward()
        scale = float(tr.state[4])
        g = torch.cat([tr.g16.float().view(-1), tr.g32]) / scale
        p_ref.grad = g.clone()
        opt.step()
        sched.step()
        tr._optimizer_step()
        torch.cuda.synchronize()
        assert int(tr.state[0].view(torch.int32)) == step + 1
        assert float(tr.g16.float().abs().max()) == 0 and float(tr.g32.abs().max()) == 0, "Adam must consume and clear the gradients"
        # eps = 1e-15 makes the update ~ lr * sign(g): entries whose gradient is ~0 are ill-conditioned, compare in norm
        assert _rel(tr.params, p_ref.detach()) < 1e-4, _rel(tr.params, p_ref.detach())
        assert torch.equal(tr.table16, tr.embeddings.half()), "fp16 shadow of the table out of date"


def test_inf_gradient_skips_step_and_halves_scale():
    tr = _make(num_rays=1024)
    tr.update_extra_state()
    tr.next_pool_batch()
    tr._forward_backward()
    tr.g32[5] = float("inf")
    before, scale0 = tr.params.clone(), float(tr.state[4])
    assert scale0 == 65536.0  # GradScaler's init_scale, as `main_nerf -O` starts (nerf/utils.py:328)
    tr._optimizer_step()
    torch.cuda.synchronize()
    assert torch.equal(before, tr.params)
    assert float(tr.state[4]) == 0.5 * scale0 and int(tr.state[0].view(torch.int32)) == 0
    assert int(tr.state[13].view(torch.int32)) == 1  # the iteration counter (learning-rate schedule) advances on a skipped step too
    assert float(tr.g32.abs().max()) == 0


@pytest.mark.parametrize("graph", [False, True])
def test_training_converges_on_the_synthetic_scene(graph):
    import nerf
    from nerf import synthetic
    tr = nerf.NGPTrainer(bound=1.0, num_rays=4096, use_graph=graph, log2_hashmap_size=17)
    o, d, gt = synthetic.make_pool(32, 4096, "cuda", seed=5)
    tr.set_pool(o, d, gt)
    losses = []
    for i in range(256):
        tr.next_pool_batch()
        tr.train_step()
        if i % 16 == 15:
            losses.append(tr.loss())
    assert np.isfinite(losses).all(), losses
    assert losses[-1] < 0.25 * losses[0], losses
    assert tr.mean_count > 0 and tr.M >= tr.mean_count


def test_graph_and_eager_steps_agree():
    """Same seed, same pool, perturb off (the only RNG draw inside a step): a captured step replays the eager sequence."""
    outs = []
    for graph in (False, True):
        torch.manual_seed(0)
        tr = _make(num_rays=1024, graph=graph, perturb=False, update_interval=0)
        torch.manual_seed(1)
        tr.update_extra_state()
        for _ in range(4):
            tr.next_pool_batch()
            tr.train_step()
        torch.cuda.synchronize()
        outs.append((tr.loss(), tr.sigma_w.clone(), tr.embeddings.clone()))
    # fp16 atomics make the table gradient order-dependent -> not bit-exact run to run
    assert abs(outs[0][0] - outs[1][0]) < 1e-3 * max(1.0, abs(outs[0][0]))
    assert _rel(outs[0][1], outs[1][1]) < 2e-2


@pytest.mark.parametrize("graph", [False, True])
def test_overlapped_optimizer_is_the_same_training(graph):
    """overlap_optimizer defers Adam(n) to run beside march(n+1): after flush() the state is that of the sequential trainer."""
    outs = []
    for overlap in (False, True):
        torch.manual_seed(0)
        tr = _make(num_rays=1024, graph=graph, perturb=False, update_interval=0, overlap_optimizer=overlap)
        torch.manual_seed(1)
        tr.update_extra_state()
        losses = []
        for _ in range(6):
            tr.next_pool_batch()
            tr.train_step()
            losses.append(tr.loss())
        if overlap:
            assert int(tr.state[0].view(torch.int32)) == 5, "one optimizer step must still be pending"
        tr.flush()
        torch.cuda.synchronize()
        assert int(tr.state[0].view(torch.int32)) == 6
        outs.append((losses, tr.sigma_w.clone(), tr.embeddings.clone(), tr.table16.clone()))
    assert np.allclose(outs[0][0], outs[1][0], rtol=1e-3, atol=1e-6), (outs[0][0], outs[1][0])
    assert _rel(outs[0][1], outs[1][1]) < 2e-2  # fp16 atomics: not bit-exact run to run (see test above)
    assert torch.equal(outs[1][3], outs[1][2].half())


def test_step_matches_cpu_oracle_step():
    """Whole step (march -> encode -> MLPs -> composite -> loss -> backward -> Adam) against oracle/nerf_step_ref.py."""
    from oracle import cpu_ref, nerf_step_ref as R
    cpu_ref.build()
    import nerf
    from nerf import synthetic
    tr = nerf.NGPTrainer(bound=1.0, num_rays=512, use_graph=False, log2_hashmap_size=15, perturb=False, update_interval=0)
    o, d, gt = synthetic.make_pool(2, 512, "cuda", seed=11)
    tr.set_pool(o, d, gt)
    tr.update_extra_state()
    st = R.RefState(tr.embeddings.cpu().numpy(), tr.sigma_w.cpu().numpy(), tr.color_w.cpu().numpy(), tr.offsets_host, tr.per_level_scale,
                    tr.base_resolution, loss_scale=float(tr.state[4]))
    bits = tr.density_bitfield.cpu().numpy()
    for step in range(2):
        tr.next_pool_batch()
        ro, rd, g = tr.rays_o.cpu().numpy(), tr.rays_d.cpu().numpy(), tr.gt.cpu().numpy()
        tr._forward_backward()
        torch.cuda.synchronize()
        loss_ref, grads, n_ref = R.forward_backward(st, ro, rd, g, bits, tr.cascade, tr.H, tr.M)
        assert int(tr.step_counter[0, 0]) == n_ref, "sample count differs from the oracle marcher"
        assert abs(float(tr.state[7]) - loss_ref) < 2e-4 * max(1.0, loss_ref), (float(tr.state[7]), loss_ref)
        gt_t = torch.from_numpy(grads[0]).cuda()
        assert _rel(tr.g16.float(), gt_t) < 2e-2, _rel(tr.g16.float(), gt_t)
        assert _rel(tr.g_sig, torch.from_numpy(grads[1]).cuda()) < 2e-2
        assert _rel(tr.g_col, torch.from_numpy(grads[2]).cuda()) < 2e-2
        # feed the device gradients to both optimizers so the parameter comparison isolates the Adam restatement
        dev_grads = (tr.g16.float().cpu().numpy(), tr.g_sig.cpu().numpy().copy(), tr.g_col.cpu().numpy().copy())
        R.adam_step(st, dev_grads)
        tr._optimizer_step()
        torch.cuda.synchronize()
        assert _rel(tr.embeddings, torch.from_numpy(st.table).cuda()) < 1e-4
        assert _rel(tr.color_w, torch.from_numpy(st.color_w).cuda()) < 1e-5


def test_device_side_batch_limit():
    """ngp_call_opts.batch_limit: the per-sample kernels stop at min(B, *count) read on the device; rows past it are
    untouched; the level-major layout keeps the caller's B as its row stride; a call made without the
    option right after is unlimited (no process-wide state)."""
    import numpy as np
    import ngp_abi as abi
    import gridencoder
    import ffmlp
    gb, fb = gridencoder._backend, ffmlp._backend
    B, n = 5000, 3217
    cnt = torch.tensor([n, 0], dtype=torch.int32, device="cuda")
    enc = gridencoder.GridEncoder(num_levels=16, level_dim=2, base_resolution=16, log2_hashmap_size=15, desired_resolution=512).cuda()
    with torch.no_grad():
        enc.embeddings.uniform_(-1, 1)
    L, S, H = 16, float(np.log2(enc.per_level_scale)), 16
    emb16, offs = enc.embeddings.detach().half().contiguous(), enc.offsets
    x = torch.rand(B, 3, device="cuda")
    net = ffmlp.FFMLP(32, 16, 64, 2).cuda()
    w = net.weights.detach()
    g = torch.randn(B, 16, device="cuda").half()

    def run(xs, gs, opts, layout, fill=7.0):
        nb = xs.shape[0]
        feat = torch.full((L, nb, 2) if layout == abi.LAYOUT_LBC else (nb, 32), fill, dtype=torch.half, device="cuda")
        gb.grid_encode_forward(xs, emb16, offs, feat, nb, 3, 2, L, S, H, None, 0, False, 0, layout=layout, opts=opts)
        rows = feat.permute(1, 0, 2).reshape(nb, 32).contiguous() if layout == abi.LAYOUT_LBC else feat
        y = torch.full((nb, 16), fill, dtype=torch.half, device="cuda")
        fb.ffmlp_forward(rows, w, nb, 32, 16, 64, 2, 0, 6, None, y, opts=opts)
        gx, gw = torch.full((nb, 32), fill, dtype=torch.half, device="cuda"), torch.empty_like(w)
        fb.ffmlp_backward(gs, rows, w, None, nb, 32, 16, 64, 2, 0, 6, True, None, gx, gw, opts=opts)
        gfeat = gx.view(nb, L, 2).permute(1, 0, 2).contiguous() if layout == abi.LAYOUT_LBC else gx
        ge = torch.zeros_like(emb16)
        gb.grid_encode_backward(gfeat, xs, emb16, offs, ge, nb, 3, 2, L, S, H, None, None, 0, False, 0, layout=layout, opts=opts)
        return feat, y, gx, gw, ge

    for layout in (abi.LAYOUT_BLC, abi.LAYOUT_LBC):
        f_ref, y_ref, gx_ref, gw_ref, ge_ref = run(x[:n].contiguous(), g[:n].contiguous(), None, layout)
        f, y, gx, gw, ge = run(x, g, abi.call_opts(batch_limit=cnt), layout)
        if layout == abi.LAYOUT_LBC:
            assert torch.equal(f[:, :n], f_ref) and bool((f[:, n:] == 7.0).all())  # stride B, rows past the limit untouched
        else:
            assert torch.equal(f[:n], f_ref) and bool((f[n:] == 7.0).all())
        assert torch.equal(y[:n], y_ref) and bool((y[n:] == 7.0).all())
        assert torch.equal(gx[:n], gx_ref) and bool((gx[n:] == 7.0).all())
        assert _rel(gw, gw_ref) < 1e-6
        assert _rel(ge, ge_ref) < 1e-2  # fp16 atomics, order-dependent
    # a limit above B is clamped to B; and the call after a limited one is unlimited
    cnt[0] = 10 * B
    _, y2, _, gw2, _ = run(x, g, abi.call_opts(batch_limit=cnt), abi.LAYOUT_BLC)
    cnt[0] = 1
    run(x, g, abi.call_opts(batch_limit=cnt), abi.LAYOUT_BLC)
    _, y_full, _, gw_full, _ = run(x, g, None, abi.LAYOUT_BLC)
    assert torch.equal(y2, y_full) and _rel(gw2, gw_full) < 1e-6 and bool((y_full != 7.0).any(dim=1).all())


def test_inference_render_matches_training_quadrature_and_band_sharding():
    """NGPTrainer.render (march_rays / composite_rays loop with device-side compaction) against the training-path
    quadrature of the same rays, and a frame rendered as two row bands against the frame rendered at once."""
    import nerf
    import raymarching
    from nerf import synthetic
    from nerf.parallel import row_band
    tr = nerf.NGPTrainer(bound=1.0, num_rays=2048, use_graph=False, log2_hashmap_size=17)
    o, d, gt = synthetic.make_pool(16, 2048, "cuda", seed=9)
    tr.set_pool(o, d, gt)
    for _ in range(96):
        tr.next_pool_batch()
        tr.train_step()
    W = H = 48
    ro, rd = synthetic.frame_rays(W, H, "cuda", pose_index=7)
    img, depth = tr.render(ro, rd)
    assert img.shape == (W * H, 3) and torch.isfinite(img).all() and torch.isfinite(depth).all()
    # the same rays through the training ops (no perturbation): identical samples, same network, same threshold
    nears, fars = raymarching.near_far_from_aabb(ro, rd, tr.aabb, tr.min_near)
    xyzs, dirs, deltas, rays = raymarching.march_rays_train(ro, rd, tr.bound, tr.density_bitfield, tr.cascade, tr.H, nears, fars, None, -1, False,
                                                             128, True, tr.dt_gamma, tr.max_steps)
    M = xyzs.shape[0]
    x01 = torch.empty(M, 3, device="cuda"); enc = torch.empty(M, 32, dtype=torch.half, device="cuda")
    h = torch.empty(M, 16, dtype=torch.half, device="cuda"); cin = torch.empty(M, 32, dtype=torch.half, device="cuda"); c = torch.empty_like(h)
    sig = torch.empty(M, device="cuda"); rgb = torch.empty(M, 3, device="cuda")
    import ngp_abi as abi
    tr._encode(xyzs.contiguous(), x01, enc, M)
    tr._sigma_forward(enc, h, M, True)
    abi.check(abi.lib.ngp_nerf_mid_forward(h.data_ptr(), dirs.contiguous().data_ptr(), tr.density_scale, sig.data_ptr(), cin.data_ptr(), M, abi.stream(), None))
    tr._color_forward(cin, c, M, True)
    abi.check(abi.lib.ngp_nerf_rgb_forward(c.data_ptr(), rgb.data_ptr(), M, abi.stream(), None))
    ws, dep, im = raymarching.composite_rays_train(sig, rgb, deltas, rays, 1e-4)
    ref = im + (1 - ws).unsqueeze(-1) * tr.bg_color
    assert (img - ref).abs().max() < 2e-3, float((img - ref).abs().max())
    # two bands == whole frame (inference shards rows over the ranks, no collective)
    parts = []
    for r in range(2):
        r0, r1 = row_band(H, r, 2)
        parts.append(tr.render(ro[r0 * W:r1 * W], rd[r0 * W:r1 * W])[0])
    assert torch.allclose(torch.cat(parts), img, atol=1e-5)
    # interleaved stripes == whole frame
    from nerf.parallel import row_stripes
    out = torch.empty_like(img)
    for r in range(3):
        rows = row_stripes(H, r, 3, stripe=4).cuda()
        idx = (rows[:, None] * W + torch.arange(W, device="cuda")[None, :]).reshape(-1)
        out[idx] = tr.render(ro[idx], rd[idx])[0]
    assert torch.allclose(out, img, atol=2e-3)
    # The loop schedule (how often the alive count is read back, how many samples per ray an iteration takes) changes
    # where a ray's march is cut and resumed from the compositor's running t, i.e. sample positions in their last bits
    # (the reference's own result depends on its n_alive-driven n_step in the same way): every schedule must agree with
    # the reference's to the tolerance of the fp16 network, not bit for bit.
    img1, depth1 = tr.render(ro, rd, sync_every=1, use_graph=False, n_step_range=(1, 8))   # the reference's schedule
    ref_iters = tr.last_render_iters
    for kw in (dict(sync_every=8), dict(sync_every=2, n_step_range=(3, 17)), dict(use_graph=False), dict(sync_every=4, n_step_range=(1, 8))):
        a, b = tr.render(ro, rd, **kw)
        assert (a - img1).abs().max() < 2e-3 and (b - depth1).abs().max() < 2e-3, (kw, float((a - img1).abs().max()), float((b - depth1).abs().max()))
    tr.render(ro, rd)
    assert tr.last_render_iters < ref_iters


def test_get_rays_matches_reference_construction():
    """nerf.get_rays (index selection in torch, ray construction in one kernel) against nerf/utils.py:69-134 transcribed."""
    import nerf
    from test_oracle import _orbit_poses
    poses = torch.from_numpy(_orbit_poses(5)).cuda()
    H, W = 60, 80
    intr = (70.0, 65.0, 39.5, 30.25)

    def ref(poses, inds):  # utils.py:69-71, 118-134
        B = poses.shape[0]
        fx, fy, cx, cy = intr
        i, j = torch.meshgrid(torch.linspace(0, W - 1, W, device="cuda"), torch.linspace(0, H - 1, H, device="cuda"), indexing="ij")
        i = i.t().reshape([1, H * W]).expand([B, H * W]) + 0.5
        j = j.t().reshape([1, H * W]).expand([B, H * W]) + 0.5
        if inds is not None:
            i, j = torch.gather(i, -1, inds), torch.gather(j, -1, inds)
        zs = torch.ones_like(i)
        d = torch.stack(((i - cx) / fx * zs, (j - cy) / fy * zs, zs), dim=-1)
        d = d / torch.norm(d, dim=-1, keepdim=True)
        rays_d = d @ poses[:, :3, :3].transpose(-1, -2)
        rays_o = poses[..., :3, 3][..., None, :].expand_as(rays_d)
        return rays_o, rays_d

    full = nerf.get_rays(poses, intr, H, W, -1)
    ro, rd = ref(poses, None)
    assert full["rays_o"].shape == (5, H * W, 3) and torch.equal(full["rays_o"], ro.contiguous())
    assert torch.allclose(full["rays_d"], rd, rtol=0, atol=2e-7)
    torch.manual_seed(3)
    some = nerf.get_rays(poses, intr, H, W, 1000)
    assert some["inds"].shape == (5, 1000) and int(some["inds"].max()) < H * W
    ro, rd = ref(poses, some["inds"])
    assert torch.allclose(some["rays_d"], rd, rtol=0, atol=2e-7) and torch.equal(some["rays_o"], ro.contiguous())
    err = torch.rand(5, 128 * 128, device="cuda")
    em = nerf.get_rays(poses, intr, H, W, 500, error_map=err)
    ro, rd = ref(poses, em["inds"])
    assert em["inds_coarse"].shape == (5, 500) and torch.allclose(em["rays_d"], rd, rtol=0, atol=2e-7)
    pt = nerf.get_rays(poses[:1], intr, H, W, 64, patch_size=4)
    assert pt["rays_d"].shape == (1, 64, 3)
    ro, rd = ref(poses[:1], pt["inds"])
    assert torch.allclose(pt["rays_d"], rd, rtol=0, atol=2e-7)


def test_checkpoint_round_trip_in_the_reference_layout(tmp_path):
    """save_checkpoint / load_checkpoint: the reference's keys, and a restored trainer continues bit-for-bit like the
    original where the step is deterministic (everything but the fp16 table-gradient atomics)."""
    import nerf
    tr = _make(num_rays=1024, graph=False, perturb=False, update_interval=0)
    tr.update_extra_state()
    for _ in range(3):
        tr.next_pool_batch(); tr.train_step()
    path = str(tmp_path / "ngp.pth")
    tr.save_checkpoint(path)
    ck = torch.load(path)
    assert {"model", "mean_count", "mean_density", "global_step", "epoch", "stats"} <= set(ck)
    # what the reference's Trainer.load_checkpoint reads from the file (nerf/utils.py:1254-1272), transcribed: every access succeeds
    ref_loader = dict(model=ck["model"], mean_count=ck["mean_count"], mean_density=ck["mean_density"], stats=ck["stats"], epoch=ck["epoch"],
                      global_step=ck["global_step"])
    assert isinstance(ref_loader["stats"], dict) and {"loss", "valid_loss", "results", "checkpoints", "best_result"} <= set(ref_loader["stats"])
    assert isinstance(ref_loader["epoch"], int) and ref_loader["global_step"] == 3 and ref_loader["mean_density"] > 0
    assert {"encoder.embeddings", "encoder.offsets", "sigma_net.weights", "color_net.weights", "density_grid", "density_bitfield",
            "step_counter", "aabb_train", "aabb_infer"} == set(ck["model"])
    tr2 = _make(num_rays=1024, graph=False, perturb=False, update_interval=0)
    tr2.load_checkpoint(path)
    assert torch.equal(tr2.params, tr.params) and torch.equal(tr2.table16, tr.table16) and torch.equal(tr2.exp_avg, tr.exp_avg)
    assert torch.equal(tr2.density_bitfield, tr.density_bitfield) and torch.equal(tr2.state, tr.state) and tr2.global_step == tr.global_step
    assert abs(tr2.mean_density() - tr.mean_density()) <= 1e-6 * tr.mean_density()  # restored, not 0
    assert torch.equal(tr2.count_sum, tr.count_sum) and tr2.local_step == tr.local_step
    # the twin module of the reference topology loads the same state dict
    net = nerf.NeRFNetwork(bound=tr.bound, log2_hashmap_size=17).cuda()
    missing, unexpected = net.load_state_dict({k: v for k, v in ck["model"].items()}, strict=False)
    assert not [k for k in missing if k.split(".")[0] in ("encoder", "sigma_net", "color_net")], missing
    ro, rd = tr._pool[0][:256], tr._pool[1][:256]
    a, _ = tr.render(ro, rd)
    b, _ = tr2.render(ro, rd)
    assert torch.equal(a, b)
