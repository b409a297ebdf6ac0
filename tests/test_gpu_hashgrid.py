"""GPU parity tests of the grid encoder (through gridencoder._backend -> C ABI -> sm_100a kernels)
against the CPU oracle, plus size-independent properties at BASELINE.json's full size.

Tolerances (north_star): fp32 1e-5 relative, fp16 1e-3 relative, of the output scale.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _be():
    import gridencoder
    return gridencoder._backend


def _abi():
    import ngp_abi
    return ngp_abi


def _run_fwd(x, emb, offsets, pls, H, calc_dy=False, gridtype=0, ac=False, interp=0, layout=0, opts=None):
    be = _be()
    B, D = x.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    out = torch.empty((L, B, C) if layout == 0 else (B, L * C), device=DEV, dtype=emb.dtype)
    dy = torch.empty(B, L * D * C, device=DEV, dtype=emb.dtype) if calc_dy else None
    be.grid_encode_forward(x, emb, offsets, out, B, D, C, L, np.log2(pls), H, dy, gridtype, ac, interp, layout=layout, opts=opts)
    return out, dy


def _run_bwd(grad, x, emb, offsets, pls, H, dy=None, gridtype=0, ac=False, interp=0, layout=0, opts=None):
    be = _be()
    B, D = x.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    g_emb = torch.zeros_like(emb)
    g_in = torch.zeros(B, D, device=DEV, dtype=emb.dtype) if dy is not None else None
    be.grid_encode_backward(grad, x, emb, offsets, g_emb, B, D, C, L, np.log2(pls), H, dy, g_in, gridtype, ac, interp, layout=layout, opts=opts)
    return g_emb, g_in


def _cfg(oracle, D, C, L=4, log2T=9, pls=2.0, H=4, ac=False, seed=0, B=777, dtype=np.float32):
    offsets, pls = oracle.make_offsets(D, L, C, pls, H, log2T, None, ac)
    rng = np.random.default_rng(seed)
    x = rng.random((B, D)).astype(np.float32)
    x[0, :] = 0.0
    x[1, :] = 1.0
    x[2, 0] = -0.25          # out of range -> zeros
    x[3, D - 1] = 1.0 + 1e-3
    emb = ((rng.random((offsets[-1], C)) - 0.5) * 2).astype(dtype)
    return offsets, pls, x, emb


@pytest.mark.parametrize("D", [2, 3, 4, 5])
@pytest.mark.parametrize("C", [1, 2, 4, 8])
def test_forward_backward_fp32_all_DC(oracle, D, C):
    H, L = 4, (3 if D >= 4 else 4)
    for gridtype, ac, interp in [(0, False, 0), (1, False, 0), (0, True, 1)]:
        offs, pls, x, embx = _cfg(oracle, D, C, L=L, ac=ac)
        ref, ref_dy = oracle.grid_encode_forward(x, embx, offs, pls, H, True, gridtype, ac, interp)
        xt, et, ot = torch.from_numpy(x).to(DEV), torch.from_numpy(embx).to(DEV), torch.from_numpy(offs).to(DEV)
        out, dy = _run_fwd(xt, et, ot, pls, H, True, gridtype, ac, interp)
        torch.cuda.synchronize()
        assert np.allclose(out.cpu().numpy(), ref, rtol=1e-5, atol=1e-5 * np.abs(ref).max())
        assert np.allclose(dy.cpu().numpy(), ref_dy, rtol=1e-4, atol=1e-4 * np.abs(ref_dy).max())
        rng = np.random.default_rng(1)
        g = rng.standard_normal(ref.shape).astype(np.float32)
        ref_ge, ref_gi = oracle.grid_encode_backward(g, x, embx, offs, pls, H, ref_dy, gridtype, ac, interp)
        ge, gi = _run_bwd(torch.from_numpy(g).to(DEV), xt, et, ot, pls, H, dy, gridtype, ac, interp)
        torch.cuda.synchronize()
        assert np.allclose(ge.cpu().numpy(), ref_ge, rtol=1e-4, atol=1e-5 * np.abs(ref_ge).max())
        assert np.allclose(gi.cpu().numpy(), ref_gi, rtol=1e-3, atol=1e-4 * np.abs(ref_gi).max())


def test_forward_backward_fp64(oracle):
    offsets, pls, x, emb = _cfg(oracle, 3, 2, dtype=np.float64)
    ref, ref_dy = oracle.grid_encode_forward(x, emb, offsets, pls, 4, True)
    xt, et, ot = torch.from_numpy(x).to(DEV), torch.from_numpy(emb).to(DEV), torch.from_numpy(offsets).to(DEV)
    out, dy = _run_fwd(xt, et, ot, pls, 4, True)
    assert np.allclose(out.cpu().numpy(), ref, rtol=1e-12, atol=1e-13)
    g = np.random.default_rng(2).standard_normal(ref.shape)
    ref_ge, ref_gi = oracle.grid_encode_backward(g, x, emb, offsets, pls, 4, ref_dy)
    ge, gi = _run_bwd(torch.from_numpy(g).to(DEV), xt, et, ot, pls, 4, dy)
    assert np.allclose(ge.cpu().numpy(), ref_ge, rtol=1e-10, atol=1e-12)
    assert np.allclose(gi.cpu().numpy(), ref_gi, rtol=1e-10, atol=1e-10)


def test_torch_gradcheck_fp64_reference_config():
    # /root/reference/testing/test_hashgrid_grad.py:24-61, run against the new kernels
    from gridencoder import grid_encode
    D, L, C, H, pls, log2T = 3, 4, 2, 4, 2, 8
    offsets, off = [], 0
    for i in range(L):
        res = int(np.ceil(H * pls ** i))
        offsets.append(off)
        off += min(2 ** log2T, (res + 1) ** D)
    offsets.append(off)
    offsets = torch.tensor(offsets, dtype=torch.int32, device=DEV)
    torch.manual_seed(0)
    inputs = torch.rand(1, D, dtype=torch.float64, device=DEV).float()
    emb = (torch.rand(off, C, dtype=torch.float64, device=DEV) * 0.1).requires_grad_(True)
    assert torch.autograd.gradcheck(lambda e: grid_encode(inputs, e, offsets, pls, H, False), (emb,), eps=1e-2, atol=1e-3, rtol=0.01,
                                    nondet_tol=1e-9)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
@pytest.mark.parametrize("layout", [0, 1])
@pytest.mark.parametrize("variant", [1, 2, 3])
def test_fast_path_equals_generic(oracle, dtype, layout, variant):
    """D=3,C=2 fast kernels (paired accesses, shared-memory coarse levels) vs the generic kernels."""
    abi = _abi()
    offsets, pls = oracle.make_offsets(desired_resolution=2048)
    B = 40000  # > 32768 so the shared-memory levels are exercised
    torch.manual_seed(0)
    x = torch.rand(B, 3, device=DEV)
    x[:7] = torch.tensor([[0, 0, 0], [1, 1, 1], [0.5, 0.5, 0.5], [-0.1, 0.2, 0.3], [0.2, 1.1, 0.3], [1, 0, 1], [0, 1, 0]], device=DEV)
    emb = ((torch.rand(int(offsets[-1]), 2, device=DEV) - 0.5) * 2).to(dtype)
    ot = torch.from_numpy(offsets).to(DEV)
    for interp, ac in [(0, False), (1, False), (0, True)]:
        offs2, _ = oracle.make_offsets(desired_resolution=2048, align_corners=ac)
        ot = torch.from_numpy(offs2).to(DEV)
        e = emb[: int(offs2[-1])].contiguous()
        abi.set_option("hashgrid_fwd_variant", 0)
        ref, _ = _run_fwd(x, e, ot, pls, 16, False, 0, ac, interp, layout)
        abi.set_option("hashgrid_fwd_variant", variant)
        out, _ = _run_fwd(x, e, ot, pls, 16, False, 0, ac, interp, layout)
        abi.set_option("hashgrid_fwd_variant", -1)
        if dtype == torch.float32:
            assert torch.equal(out, ref)  # same operation order -> bit-identical
        else:
            assert torch.allclose(out.float(), ref.float(), rtol=2e-3, atol=2e-3)
        g = torch.randn_like(ref)
        abi.set_option("hashgrid_bwd_variant", 0)
        ge_ref, _ = _run_bwd(g, x, e, ot, pls, 16, None, 0, ac, interp, layout)
        abi.set_option("hashgrid_bwd_variant", variant)
        for agg in (0, 1):
            abi.set_option("hashgrid_bwd_warp_agg", agg)
            ge, _ = _run_bwd(g, x, e, ot, pls, 16, None, 0, ac, interp, layout)
            if dtype == torch.float32:
                assert torch.allclose(ge, ge_ref, rtol=1e-4, atol=1e-4)
            else:
                # coarse levels: thousands of fp16 atomic adds per cell in the generic path vs one fp32 sum here
                hashed = slice(int(offs2[6]), None)
                assert torch.allclose(ge[hashed].float(), ge_ref[hashed].float(), rtol=2e-2, atol=2e-2)
        abi.set_option("hashgrid_bwd_warp_agg", 1)
        abi.set_option("hashgrid_bwd_variant", -1)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
@pytest.mark.parametrize("gridtype", [0, 1])
def test_v3_kernel_options_equal_generic(oracle, dtype, gridtype):
    """Every switch of the per-level-kind kernels (warp-mixed level groups, occupancy target, 16 B group reductions,
    zero-gradient skip) and its general-index fallback (tiled grids: levels that wrap) against the generic kernels;
    random and ray-ordered points, a block of exactly zero upstream gradients."""
    abi = _abi()
    offsets, pls = oracle.make_offsets(desired_resolution=2048)
    ot = torch.from_numpy(offsets).to(DEV)
    B = 40000  # > 32768: the shared-memory coarse levels of NGP_OPT_GRID_BWD_SMEM_COARSE are exercised
    torch.manual_seed(3)
    x = torch.rand(B, 3, device=DEV)
    n_rays = 64
    o = torch.rand(n_rays, 1, 3, device=DEV) * 0.5 + 0.25
    d = torch.nn.functional.normalize(torch.randn(n_rays, 1, 3, device=DEV), dim=-1)
    t = torch.arange(128, device=DEV).view(1, 128, 1) * (3.4641 / 2048)
    x[: n_rays * 128] = (o + d * t).clamp(0, 1).reshape(-1, 3)  # coherent part: runs of equal cells
    x[-3:] = torch.tensor([[-0.1, 0.2, 0.3], [0, 0, 0], [1, 1, 1]], device=DEV)
    emb = ((torch.rand(int(offsets[-1]), 2, device=DEV) - 0.5) * 2).to(dtype)
    g = torch.randn(B, 32, device=DEV).to(dtype)
    g[100:400] = 0  # zero-gradient lanes are skipped
    keys = ("hashgrid_fwd_variant", "hashgrid_bwd_variant", "hashgrid_v3_mix_fwd", "hashgrid_v3_mix_bwd", "hashgrid_v3_minb_fwd",
            "hashgrid_v3_minb_bwd", "hashgrid_bwd_red4")
    defaults = dict(zip(keys, (-1, -1, 0, 1, -1, -1, 2)))

    def setopts(**kw):
        for k, v in {**defaults, **kw}.items():
            abi.set_option(k, v)
    try:
        setopts(hashgrid_fwd_variant=0, hashgrid_bwd_variant=0)
        ref, _ = _run_fwd(x, emb, ot, pls, 16, False, gridtype, False, 0, 1)
        ge_ref, _ = _run_bwd(g, x, emb, ot, pls, 16, None, gridtype, False, 0, 1)
        for kw in (dict(), dict(hashgrid_v3_mix_fwd=1, hashgrid_v3_mix_bwd=0), dict(hashgrid_v3_minb_fwd=4, hashgrid_v3_minb_bwd=6),
                   dict(hashgrid_v3_minb_fwd=6, hashgrid_bwd_red4=0), dict(hashgrid_bwd_red4=1), dict(coarse_flag=1)):
            kw = dict(kw)
            coarse = abi.call_opts(flags=abi.OPT_GRID_BWD_SMEM_COARSE) if kw.pop("coarse_flag", 0) else None  # per-call flag, not an option
            setopts(hashgrid_fwd_variant=3, hashgrid_bwd_variant=3, **kw)
            out, _ = _run_fwd(x, emb, ot, pls, 16, False, gridtype, False, 0, 1)
            ge, _ = _run_bwd(g, x, emb, ot, pls, 16, None, gridtype, False, 0, 1, opts=coarse)
            if dtype == torch.float32:
                assert torch.equal(out, ref), kw
                assert torch.allclose(ge, ge_ref, rtol=1e-4, atol=1e-4), kw
            else:
                assert torch.allclose(out.float(), ref.float(), rtol=2e-3, atol=2e-3), kw
                hashed = slice(int(offsets[6]), None)  # coarse levels: fp16 atomic order noise of the generic path
                assert torch.allclose(ge[hashed].float(), ge_ref[hashed].float(), rtol=2e-2, atol=2e-2), kw
    finally:
        setopts()


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
def test_input_bound_is_the_module_mapping(oracle, dtype):
    """ngp_call_opts.input_bound = b: the kernels map [-b, b] -> [0, 1] as they load the points, bit for bit the
    (x + b) / (2 b) of GridEncoder.forward (grid.py:807); unsupported configurations refuse loudly.  The option travels
    with the call: a call without it, made right after, sees inputs in [0, 1] again (no process-wide state)."""
    abi = _abi()
    offsets, pls = oracle.make_offsets(desired_resolution=2048)
    ot = torch.from_numpy(offsets).to(DEV)
    B, bound = 50000, 2.0
    torch.manual_seed(5)
    x = (torch.rand(B, 3, device=DEV) * 2 - 1) * bound * 1.01  # a few points outside
    x01 = (x + bound) / (2 * bound)
    emb = ((torch.rand(int(offsets[-1]), 2, device=DEV) - 0.5) * 2).to(dtype)
    g = torch.randn(B, 32, device=DEV).to(dtype)
    ref, _ = _run_fwd(x01, emb, ot, pls, 16, layout=1)
    for coarse in (0, abi.OPT_GRID_BWD_SMEM_COARSE):
        ge_ref, _ = _run_bwd(g, x01, emb, ot, pls, 16, layout=1, opts=abi.call_opts(flags=coarse))
        o = abi.call_opts(input_bound=bound, flags=coarse)
        out, _ = _run_fwd(x, emb, ot, pls, 16, layout=1, opts=o)
        ge, _ = _run_bwd(g, x, emb, ot, pls, 16, layout=1, opts=o)
        with pytest.raises(RuntimeError):  # generic kernels (no host offsets / D != 3) do not take the mapping
            _run_fwd(x[:, :2].contiguous(), emb, ot, pls, 16, layout=1, opts=o)
        again, _ = _run_fwd(x01, emb, ot, pls, 16, layout=1)  # nothing lingers after a call that carried the option
        assert torch.equal(again, ref)
        assert torch.equal(out, ref)
        if dtype == torch.float32:
            assert torch.allclose(ge, ge_ref, rtol=1e-4, atol=1e-4)
        else:
            hashed = slice(int(offsets[6]), None)
            assert torch.allclose(ge[hashed].float(), ge_ref[hashed].float(), rtol=2e-2, atol=2e-2)


def test_exact_input_grad_is_opt_in(oracle):
    """Default dy_dx = the reference's (`pos_deriv[D] = {1.0f}`, gridencoder.cu:143: zero for d > 0 with linear
    interpolation); NGP_OPT_EXACT_INPUT_GRAD / GridEncoder(exact_input_grad=True) gives the exact derivative.  Both
    against the oracle in the matching mode."""
    abi = _abi()
    offs, pls, x, emb = _cfg(oracle, 3, 2, L=4)
    xt, et, ot = torch.from_numpy(x).to(DEV), torch.from_numpy(emb).to(DEV), torch.from_numpy(offs).to(DEV)
    ref_out, ref_dy = oracle.grid_encode_forward(x, emb, offs, pls, 4, True)
    out, dy = _run_fwd(xt, et, ot, pls, 4, True)
    assert np.allclose(dy.cpu().numpy(), ref_dy, rtol=1e-4, atol=1e-4 * np.abs(ref_dy).max())
    d3 = dy.view(x.shape[0], 4, 3, 2)
    assert float(d3[:, :, 1:].abs().max()) == 0.0 and float(d3[:, :, 0].abs().max()) > 0.0
    oracle.set_pos_deriv_quirk(False)
    try:
        _, ex_dy = oracle.grid_encode_forward(x, emb, offs, pls, 4, True)
    finally:
        oracle.set_pos_deriv_quirk(True)
    out2, dy2 = _run_fwd(xt, et, ot, pls, 4, True, opts=abi.call_opts(flags=abi.OPT_EXACT_INPUT_GRAD))
    assert torch.equal(out2, out)
    assert np.allclose(dy2.cpu().numpy(), ex_dy, rtol=1e-4, atol=1e-4 * np.abs(ex_dy).max())
    assert float(dy2.view(x.shape[0], 4, 3, 2)[:, :, 1:].abs().max()) > 0.0
    # module level
    from gridencoder import GridEncoder
    torch.manual_seed(0)
    grads = {}
    for exact in (False, True):
        enc = GridEncoder(num_levels=4, desired_resolution=64, log2_hashmap_size=12, exact_input_grad=exact).to(DEV)
        with torch.no_grad():
            enc.embeddings.copy_(torch.linspace(-1, 1, enc.embeddings.numel(), device=DEV).view_as(enc.embeddings).sin())
        xr = (torch.from_numpy(x[4:]).to(DEV) * 2 - 1).requires_grad_(True)
        enc(xr).pow(2).sum().backward()
        grads[exact] = xr.grad.clone()
    assert float(grads[False][:, 1:].abs().max()) == 0.0 and float(grads[True][:, 1:].abs().max()) > 0.0
    assert torch.allclose(grads[False][:, 0], grads[True][:, 0], rtol=1e-5, atol=1e-6)


def test_warp_aggregated_backward_coherent_points(oracle):
    """Samples marching along rays: many warps sit in one coarse cell -> the shuffle-reduced path."""
    offsets, pls = oracle.make_offsets(desired_resolution=2048)
    B = 8192  # < 32768: no shared-memory levels, every level goes through the L2 reduction kernel
    t = np.linspace(0, 1, B, dtype=np.float32)
    x = np.stack([0.1 + 0.8 * t, 0.3 + 1e-4 * np.arange(B) % 0.01, np.full(B, 0.52, np.float32)], -1).astype(np.float32)
    rng = np.random.default_rng(0)
    emb = rng.standard_normal((offsets[-1], 2)).astype(np.float32)
    g = rng.standard_normal((16, B, 2)).astype(np.float32)
    ref_ge, _ = oracle.grid_encode_backward(g, x, emb, offsets, pls, 16)
    ge, _ = _run_bwd(torch.from_numpy(g).to(DEV), torch.from_numpy(x).to(DEV), torch.from_numpy(emb).to(DEV),
                     torch.from_numpy(offsets).to(DEV), pls, 16)
    assert np.allclose(ge.cpu().numpy(), ref_ge, rtol=1e-4, atol=1e-4 * np.abs(ref_ge).max())


def test_config1_slice_vs_oracle_fp32_fp16(oracle, request):
    """BASELINE config 1 geometry (L=16 F=2 T=2^19, desired_resolution 2048) on 2^14 points vs the oracle."""
    offsets, pls = oracle.make_offsets(desired_resolution=2048)
    # the level scale uses the device's exp2f; give the oracle the same 16 values (see ngp_oracle.c)
    S32 = torch.tensor(np.float32(np.log2(pls)), device=DEV)
    oracle.set_level_exp2(torch.exp2(torch.arange(16, device=DEV, dtype=torch.float32) * S32).cpu().numpy())
    request.addfinalizer(lambda: oracle.set_level_exp2(None))
    rng = np.random.default_rng(0)
    B = 1 << 14
    x = rng.random((B, 3)).astype(np.float32)
    emb = ((rng.random((offsets[-1], 2)) - 0.5) * 2).astype(np.float32)
    ref, _ = oracle.grid_encode_forward(x, emb, offsets, pls, 16)
    xt, ot = torch.from_numpy(x).to(DEV), torch.from_numpy(offsets).to(DEV)
    out, _ = _run_fwd(xt, torch.from_numpy(emb).to(DEV), ot, pls, 16)
    got = out.cpu().numpy()
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-5 * np.abs(ref).max())
    out16, _ = _run_fwd(xt, torch.from_numpy(emb).to(DEV).half(), ot, pls, 16)
    ref16, _ = oracle.grid_encode_forward(x, emb.astype(np.float16).astype(np.float32), offsets, pls, 16)
    assert np.allclose(out16.float().cpu().numpy(), ref16, rtol=1e-3, atol=1e-3 * np.abs(ref16).max())
    g = rng.standard_normal(ref.shape).astype(np.float32)
    ref_ge, _ = oracle.grid_encode_backward(g, x, emb, offsets, pls, 16)
    ge, _ = _run_bwd(torch.from_numpy(g).to(DEV), xt, torch.from_numpy(emb).to(DEV), ot, pls, 16)
    assert np.allclose(ge.cpu().numpy(), ref_ge, rtol=1e-4, atol=1e-5 * np.abs(ref_ge).max())


def test_full_size_properties_config1():
    """2^18 points, L=16, F=2, T=2^19: linearity in the table and <g, fwd(E)> == <bwd(g), E> (exact adjoint)."""
    from gridencoder import GridEncoder
    torch.manual_seed(0)
    enc = GridEncoder(desired_resolution=2048).to(DEV)
    B = 1 << 18
    x = torch.rand(B, 3, device=DEV) * 2 - 1
    E1 = (torch.rand_like(enc.embeddings) - 0.5) * 2
    E2 = (torch.rand_like(enc.embeddings) - 0.5) * 2
    with torch.no_grad():
        enc.embeddings.copy_(E1); y1 = enc(x)
        enc.embeddings.copy_(E2); y2 = enc(x)
        enc.embeddings.copy_(0.5 * E1 - 2.0 * E2); y3 = enc(x)
    assert y1.shape == (B, 32) and torch.isfinite(y1).all()
    assert torch.allclose(y3, 0.5 * y1 - 2.0 * y2, rtol=1e-4, atol=1e-5)
    enc.embeddings.data.copy_(E1)
    g = torch.randn(B, 32, device=DEV)
    y = enc(x)
    y.backward(g)
    lhs = (y.double() * g.double()).sum()
    rhs = (enc.embeddings.grad.double() * E1.double()).sum()
    assert torch.isclose(lhs, rhs, rtol=1e-5), (lhs.item(), rhs.item())
    # every point touches <= 8 entries per level; untouched entries keep an exactly-zero gradient
    assert (enc.embeddings.grad != 0).any(dim=1).sum().item() <= 8 * 16 * B


def test_module_autocast_and_input_grad():
    from gridencoder import GridEncoder
    torch.manual_seed(0)
    enc = GridEncoder(num_levels=8, desired_resolution=256, log2_hashmap_size=14).to(DEV)
    with torch.no_grad():
        enc.embeddings.uniform_(-1, 1)
    x = (torch.rand(2, 500, 3, device=DEV) * 2 - 1)
    y32 = enc(x, bound=1)
    assert y32.dtype == torch.float32 and y32.shape == (2, 500, 16)
    with torch.autocast("cuda", dtype=torch.float16):
        y16 = enc(x, bound=1)
    assert y16.dtype == torch.float16
    assert torch.allclose(y16.float(), y32, rtol=2e-3, atol=2e-3)
    with torch.autocast("cuda", dtype=torch.float16):
        enc(x).float().pow(2).sum().backward()
    assert enc.embeddings.grad is not None and enc.embeddings.grad.dtype == torch.float32  # autograd casts back to the parameter dtype
    xr = x.clone().requires_grad_(True)
    enc(xr).pow(2).sum().backward()
    assert xr.grad is not None and xr.grad.shape == x.shape and xr.grad.abs().sum() > 0


def test_grad_total_variation(oracle):
    offsets, pls, x, emb = _cfg(oracle, 3, 2, L=4, log2T=9, B=300)
    be = _be()
    grad = np.zeros_like(emb)
    ref = oracle.grad_total_variation(x, emb, grad.copy(), offsets, 0.37, pls, 4)
    gt = torch.zeros(emb.shape, device=DEV)
    be.grad_total_variation(torch.from_numpy(x).to(DEV), torch.from_numpy(emb).to(DEV), gt, torch.from_numpy(offsets).to(DEV), 0.37,
                            x.shape[0], 3, 2, 4, np.log2(pls), 4, 0, False)
    assert np.allclose(gt.cpu().numpy(), ref, rtol=1e-4, atol=1e-6)


def test_error_behaviour():
    be = _be()
    x = torch.rand(4, 3)
    e = torch.rand(64, 2)
    o = torch.tensor([0, 64], dtype=torch.int32)
    out = torch.empty(1, 4, 2)
    with pytest.raises(RuntimeError, match="must be a CUDA tensor"):
        be.grid_encode_forward(x, e, o, out, 4, 3, 2, 1, 1.0, 4, None, 0, False, 0)
    xc, oc = x.to(DEV), o.to(DEV)
    with pytest.raises(RuntimeError, match="must be a contiguous tensor"):
        be.grid_encode_forward(xc, torch.rand(2, 64, device=DEV).t(), oc, out.to(DEV), 4, 3, 2, 1, 1.0, 4, None, 0, False, 0)
    e3 = torch.rand(64, 3, device=DEV)
    with pytest.raises(RuntimeError, match="C must be 1, 2, 4, or 8"):
        be.grid_encode_forward(xc, e3, oc, torch.empty(1, 4, 3, device=DEV), 4, 3, 3, 1, 1.0, 4, None, 0, False, 0)
    with pytest.raises(RuntimeError, match="must be an int tensor"):
        be.grid_encode_forward(xc, e.to(DEV), oc.long(), out.to(DEV), 4, 3, 2, 1, 1.0, 4, None, 0, False, 0)
    # empty batch is a no-op
    be.grid_encode_forward(torch.empty(0, 3, device=DEV), e.to(DEV), oc, torch.empty(1, 0, 2, device=DEV), 0, 3, 2, 1, 1.0, 4, None, 0, False, 0)


@pytest.mark.parametrize("layout", [0, 1])
def test_segmented_warp_reduction_short_rays(oracle, layout):
    """Runs of equal cells that start and end at arbitrary lanes, with out-of-range points in between (segmented scan)."""
    offsets, pls = oracle.make_offsets(desired_resolution=2048)
    rng = np.random.default_rng(3)
    segs = []
    for i in range(300):
        n = int(rng.integers(1, 90))
        o, d = rng.random(3) * 0.6 + 0.2, rng.standard_normal(3)
        d /= np.linalg.norm(d)
        p = o + d * (np.arange(n)[:, None] * 0.0034)
        segs.append(p)
        if i % 7 == 0:
            segs.append(np.array([[-0.2, 0.5, 0.5], [0.5, 1.3, 0.5]]))  # skipped by the kernels (outside [0,1])
    x = np.concatenate(segs).astype(np.float32)
    B = x.shape[0]
    emb = rng.standard_normal((offsets[-1], 2)).astype(np.float32)
    g = rng.standard_normal((16, B, 2)).astype(np.float32)
    ref_ge, _ = oracle.grid_encode_backward(g, x, emb, offsets, pls, 16)
    g_t = torch.from_numpy(g).to(DEV)
    if layout == 1:
        g_t = g_t.permute(1, 0, 2).contiguous()
    ge, _ = _run_bwd(g_t, torch.from_numpy(x).to(DEV), torch.from_numpy(emb).to(DEV), torch.from_numpy(offsets).to(DEV), pls, 16, layout=layout)
    assert np.allclose(ge.cpu().numpy(), ref_ge, rtol=1e-4, atol=1e-4 * np.abs(ref_ge).max())
