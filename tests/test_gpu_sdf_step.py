"""SDF training step (nerf/sdf_trainer.py, configs[4]) against the autograd path through the plugin modules
(GridEncoder + FFMLP + mape loss, i.e. sdf/netowrk_ff.py + sdf/utils.py::train_step) and against the CPU oracle step."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a.double() - b.double()).norm() / max(1e-30, b.double().norm()))


def _make(B=1 << 14, graph=False):
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    tr = SDFTrainer(batch=B, use_graph=graph, log2_hashmap_size=15, loss_scale=1024.0)
    p, y = sample_sdf_batch(B, "cuda", seed=1)
    return tr, p, y


def test_sdf_step_matches_autograd_path_and_oracle(oracle):
    import gridencoder, ffmlp
    from oracle import nerf_step_ref as R
    tr, p, y = _make()
    tr.points.copy_(p); tr.sdfs.copy_(y)
    enc = gridencoder.GridEncoder(num_levels=16, log2_hashmap_size=15, desired_resolution=2048).cuda()
    net = ffmlp.FFMLP(32, 1, 64, 3).cuda().train()
    with torch.no_grad():
        enc.embeddings.copy_(tr.embeddings); net.weights.copy_(tr.mlp_w)
    st = R.SDFRefState(tr.embeddings.cpu().numpy(), tr.mlp_w.cpu().numpy(), tr.offsets_host, tr.per_level_scale, tr.base_resolution, loss_scale=1024.0)
    tr._forward_backward()
    torch.cuda.synchronize()
    with torch.autocast("cuda", dtype=torch.half):
        pred = net(enc(p, bound=1))
    diff = (pred.float().squeeze(-1) - y).abs() / (y.abs() + 1e-2)
    loss = diff.mean()
    (loss * 1024.0).backward()
    assert abs(float(tr.state[7]) - float(loss.detach())) < 1e-3 * max(1.0, float(loss.detach()))
    assert _rel(tr.g32, net.weights.grad) < 2e-2
    assert _rel(tr.g16.float(), enc.embeddings.grad) < 2e-2
    loss_ref, grads = R.sdf_forward_backward(st, p.cpu().numpy(), y.cpu().numpy())
    assert abs(float(tr.state[7]) - loss_ref) < 1e-3 * max(1.0, loss_ref)
    assert _rel(tr.g32, torch.from_numpy(grads[1]).cuda()) < 2e-2
    assert _rel(tr.g16.float(), torch.from_numpy(grads[0]).cuda()) < 2e-2
    # optimizer incl. the MLP weight decay, fed with the device gradients
    R.sdf_adam_step(st, (tr.g16.float().cpu().numpy(), tr.g32.cpu().numpy().copy()))
    tr._optimizer_step()
    torch.cuda.synchronize()
    assert _rel(tr.mlp_w, torch.from_numpy(st.mlp_w).cuda()) < 1e-5
    assert _rel(tr.embeddings, torch.from_numpy(st.table).cuda()) < 1e-4


@pytest.mark.parametrize("graph", [False, True])
def test_sdf_training_converges(graph):
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    tr = SDFTrainer(batch=1 << 16, use_graph=graph, log2_hashmap_size=17, lr=1e-3)
    batches = [sample_sdf_batch(1 << 16, "cuda", seed=s) for s in range(4)]
    losses = []
    for i in range(200):
        tr.train_step(*batches[i % 4])
        if i % 20 == 19:
            losses.append(tr.loss())
    assert np.isfinite(losses).all() and losses[-1] < 0.5 * losses[0], losses
    p, y = sample_sdf_batch(1 << 14, "cuda", seed=99)
    err = (tr.predict(p) - y).abs().mean()
    assert float(err) < 0.05, float(err)


@pytest.mark.parametrize("dtype", [torch.float32, torch.half])
@pytest.mark.parametrize("coarse", [0, 1])
def test_grid_backward_by_level_ranges_equals_one_call(dtype, coarse):
    """ngp_call_opts.level_begin / level_end: the table gradient accumulated range by range (finest first, as the SDF
    pipeline issues it) is the gradient of one call over every level, up to the order of the atomic additions."""
    import ngp_abi as abi
    import gridencoder
    B, L = 50000, 16
    enc = gridencoder.GridEncoder(num_levels=L, log2_hashmap_size=15, desired_resolution=2048)
    off_h = np.ascontiguousarray(enc.offsets.numpy().astype(np.int32))
    off = enc.offsets.cuda()
    T = int(off_h[-1])
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(B, 3, generator=g) * 2 - 1).cuda()
    grad = torch.randn(B, 2 * L, generator=g).cuda().to(dtype).contiguous()
    code = abi.DTYPE_F32 if dtype == torch.float32 else abi.DTYPE_F16
    flags = abi.OPT_GRID_BWD_SMEM_COARSE if coarse else 0
    S = float(np.log2(enc.per_level_scale))

    def run(out, **kw):
        return abi.lib.ngp_grid_encode_backward(grad.data_ptr(), x.data_ptr(), None, off.data_ptr(), off_h.ctypes.data, out.data_ptr(), B, 3, 2, L, S,
                                                enc.base_resolution, None, None, 0, 0, 0, code, abi.LAYOUT_BLC, abi.stream(),
                                                abi.call_opts(input_bound=1.0, flags=flags, **kw))
    whole, parts = torch.zeros(T, 2, dtype=dtype, device="cuda"), torch.zeros(T, 2, dtype=dtype, device="cuda")
    assert run(whole) == 0
    for lb, le in [(12, 16), (4, 12), (0, 4)]:
        before = parts.clone()
        assert run(parts, level_begin=lb, level_end=le) == 0, abi.lib.ngp_last_error().decode()
        touched = (parts != before).any(dim=1).nonzero().flatten()
        assert touched.numel() > 0 and int(touched.min()) >= off_h[lb] and int(touched.max()) < off_h[le]  # nothing outside the range
    torch.cuda.synchronize()
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert _rel(parts.float(), whole.float()) < tol
    assert run(parts, level_begin=2, level_end=8) != 0 and "multiples of 4" in abi.lib.ngp_last_error().decode()
    assert run(parts, level_begin=8, level_end=20) != 0


@pytest.mark.parametrize("graph", [False, True])
def test_sdf_pipelined_table_gradient_equals_the_one_call_step(graph):
    """SDFTrainer(grad_chunks=4) -- the multi-GPU schedule (grid backward by level range, collective of a finished range on
    a side stream), here with one rank so the collectives are no-ops -- trains like grad_chunks=1."""
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    B = 1 << 15
    a = SDFTrainer(batch=B, use_graph=graph, log2_hashmap_size=15, lr=1e-3, grad_chunks=1)
    b = SDFTrainer(batch=B, use_graph=graph, log2_hashmap_size=15, lr=1e-3, grad_chunks=4)
    assert b._level_ranges() == [(12, 16), (8, 12), (4, 8), (0, 4)] and a._level_ranges() == [(0, 16)]
    batches = [sample_sdf_batch(B, "cuda", seed=s) for s in range(3)]
    la, lb = [], []
    for i in range(12):
        a.train_step(*batches[i % 3]); b.train_step(*batches[i % 3])
        la.append(a.loss()); lb.append(b.loss())
        if i == 0:
            assert _rel(b.g16.float(), a.g16.float()) < 2e-2 and _rel(b.g32, a.g32) < 1e-3
    assert np.isfinite(lb).all() and np.allclose(la, lb, rtol=2e-2), (la, lb)
    a.flush(); b.flush()
    # (the table itself is not compared: Adam turns the sign of a rounding-level gradient into a full +-lr update)
    assert _rel(b.mlp_w, a.mlp_w) < 2e-2
    assert lb[-1] < lb[0]


@pytest.mark.parametrize("bits", [5, 6, 7])
def test_bin_points_is_a_morton_ordered_permutation(bits):
    """ngp_bin_points: the output is the input batch, payload attached, in non-decreasing Morton order of the (2^bits)^3 grid;
    the histogram part of the workspace is zero again afterwards (the next call starts from it)."""
    import ngp_abi as abi
    from nerf.sdf_trainer import sample_sdf_batch
    B = 100003
    p, y = sample_sdf_batch(B, "cuda", seed=5)
    p[17] = float("nan")  # lands in cell 0 of its axis, does not crash
    p[:, 0][100:200] = 1.0  # the upper face belongs to the last cell
    ws = torch.zeros(int(abi.lib.ngp_bin_points_workspace(B, bits)), dtype=torch.uint8, device="cuda")
    po, yo = torch.empty_like(p), torch.empty_like(y)
    for _ in range(2):  # the second call relies on the zeroed histogram
        po.fill_(-7.0)
        assert abi.lib.ngp_bin_points(p.data_ptr(), y.data_ptr(), B, 1.0, bits, ws.data_ptr(), ws.numel(), po.data_ptr(), yo.data_ptr(), abi.stream()) == 0
        torch.cuda.synchronize()
        assert int(ws[:4 << (3 * bits)].view(torch.int32).abs().sum()) == 0
        R = 1 << bits
        c = ((po.nan_to_num(nan=-1.0) + 1) * 0.5).clamp(0, 1).mul(R).long().clamp(max=R - 1)
        key = torch.zeros(B, dtype=torch.long, device="cuda")
        for bit in range(bits):
            for d in range(3):
                key |= ((c[:, d] >> bit) & 1) << (3 * bit + d)
        assert bool((key[1:] >= key[:-1]).all())

        def rows(a, b):  # sortable view of (point, payload) rows; NaN -> a fixed number
            m = torch.cat([a.nan_to_num(nan=123.0), b[:, None]], 1).cpu().numpy()
            return m[np.lexsort(m.T[::-1])]
        assert np.array_equal(rows(p, y), rows(po, yo))
    assert abi.lib.ngp_bin_points(p.data_ptr(), None, B, 1.0, 4, ws.data_ptr(), ws.numel(), po.data_ptr(), None, abi.stream()) != 0


def test_sdf_binned_step_equals_the_unordered_step():
    """bin_bits only changes the ORDER of the batch inside the step: loss and gradients are those of the unordered step."""
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    B = 1 << 16
    a = SDFTrainer(batch=B, use_graph=False, log2_hashmap_size=15, bin_bits=0)
    b = SDFTrainer(batch=B, use_graph=False, log2_hashmap_size=15, bin_bits=7)
    p, y = sample_sdf_batch(B, "cuda", seed=2)
    for t in (a, b):
        t.points.copy_(p); t.sdfs.copy_(y)
        t._forward_backward()
    torch.cuda.synchronize()
    assert abs(float(a.state[7]) - float(b.state[7])) < 1e-4 * max(1.0, abs(float(a.state[7])))
    assert _rel(b.g32, a.g32) < 1e-3 and _rel(b.g16.float(), a.g16.float()) < 2e-2


@pytest.mark.parametrize("graph", [False, True])
def test_sdf_deferred_optimizer_equals_the_immediate_one(graph):
    """overlap_optimizer: Adam of step n runs on a side stream beside the binning of batch n+1 -- same updates, same order."""
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    B = 1 << 15
    a = SDFTrainer(batch=B, use_graph=graph, log2_hashmap_size=15, lr=1e-3, overlap_optimizer=False)
    b = SDFTrainer(batch=B, use_graph=graph, log2_hashmap_size=15, lr=1e-3, overlap_optimizer=True)
    batches = [sample_sdf_batch(B, "cuda", seed=s) for s in range(3)]
    la, lb = [], []
    for i in range(10):
        a.train_step(*batches[i % 3]); b.train_step(*batches[i % 3])
        la.append(a.loss()); lb.append(b.loss())
    assert b._pending and not a._pending
    b.flush()
    assert not b._pending and abs(b.loss() - lb[-1]) < 1e-6  # the value moved from the accumulator to the last-loss slot
    assert np.isfinite(lb).all() and np.allclose(la, lb, rtol=2e-2), (la, lb)
    assert _rel(b.mlp_w, a.mlp_w) < 2e-2
    assert 0 < int(b.state[0:1].view(torch.int32)) == int(a.state[0:1].view(torch.int32)) <= 10  # optimizer steps taken (overflow skips included)


@pytest.mark.parametrize("B", [1 << 14, 70001])
def test_sdf_forward_and_loss_in_one_launch_equal_the_two_kernels(B):
    """ngp_sdf_forward_loss: the gradient row it writes is bit-identical to ngp_ffmlp_forward + ngp_sdf_loss, the loss agrees up to
    the order of its additions, and so does everything downstream."""
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    a = SDFTrainer(batch=B, use_graph=False, log2_hashmap_size=15, fused_loss=False, bin_bits=0)
    b = SDFTrainer(batch=B, use_graph=False, log2_hashmap_size=15, fused_loss=True, bin_bits=0)
    assert b.fused_loss and not a.fused_loss
    p, y = sample_sdf_batch(B, "cuda", seed=4)
    for t in (a, b):
        t.points.copy_(p); t.sdfs.copy_(y)
        t.g_h.fill_(7.0)
        t._forward_backward()
    torch.cuda.synchronize()
    assert torch.equal(a.g_h.view(torch.int16), b.g_h.view(torch.int16))
    la, lb = float(a.state[7]), float(b.state[7])
    assert abs(la - lb) < 1e-5 * max(1.0, abs(la)) and lb > 0
    assert _rel(b.g32, a.g32) < 1e-3 and _rel(b.g16.float(), a.g16.float()) < 2e-2
