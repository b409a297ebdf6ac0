"""Golden vectors produced by the reference's own CUDA extensions on a B200 (tests/golden/make_golden.py):
  * CPU part (no GPU): pins the oracle -- oracle/cpu_ref.py must reproduce them (bit-exact for marching / packbits /
    morton / counts, 1e-5 fp32, 1e-3 fp16 for encodings, composites and gradients);
  * GPU part (-m gpu): the new kernels, through the plugin API / C ABI, against the same vectors.
The reference has no golden files of its own for this path; these are its outputs run here."""
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")


def load(name):
    path = os.path.join(GOLD, name + ".npz")
    if not os.path.exists(path):
        pytest.skip(f"{path} not generated")
    return np.load(path)


def rel(a, b):
    return float(np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64)).max() / max(1e-30, np.abs(b).max()))


# ---------------------------------------------------------------------------------------------- oracle vs golden (CPU)
@pytest.mark.parametrize("tag,tol", [("f32", 1e-5), ("f16", 2e-3)])
def test_oracle_grid_vs_golden(oracle, tag, tol):
    g = load("grid_" + tag)
    oracle.set_level_exp2(g["level_exp2"])  # device exp2f of the per-level scale, recorded with the fixture
    try:
        out, dy = oracle.grid_encode_forward(g["x"], g["embeddings"], g["offsets"], float(g["per_level_scale"]), int(g["base_resolution"]), True)
        ge, gi = oracle.grid_encode_backward(g["grad"], g["x"], g["embeddings"], g["offsets"], float(g["per_level_scale"]), int(g["base_resolution"]), dy)
    finally:
        oracle.set_level_exp2(None)
    assert rel(out, g["outputs"]) < tol
    assert rel(dy, g["dy_dx"]) < tol * 4
    assert rel(ge, g["grad_embeddings"]) < tol * 4
    assert rel(gi, g["grad_inputs"]) < max(tol * 10, 1e-4)


def test_oracle_raymarching_utils_vs_golden(oracle):
    g = load("raymarching_utils")
    n, f = oracle.near_far_from_aabb(g["rays_o"], g["rays_d"], g["aabb"], float(g["min_near"]))
    assert np.array_equal(n, g["nears"]) and np.array_equal(f, g["fars"])
    assert np.allclose(oracle.sph_from_ray(g["rays_o"], g["rays_d"], float(g["radius"])), g["sph"], atol=1e-6)
    assert np.array_equal(oracle.morton3D(g["coords"]), g["morton"])
    assert np.array_equal(oracle.morton3D_invert(g["morton"]), g["coords_back"])
    assert np.array_equal(oracle.packbits(g["grid"], float(g["thresh"])), g["bitfield"])


@pytest.mark.parametrize("tag", ["g0", "g128"])
def test_oracle_march_and_composite_train_vs_golden(oracle, tag):
    g = load("march_train_" + tag)
    N = g["rays_o"].shape[0]
    M = N * 256
    x, d, l, rays, counter = oracle.march_rays_train(g["rays_o"], g["rays_d"], float(g["bound"]), g["bitfield"], int(g["C"]), int(g["H"]), g["nears"],
                                                     g["fars"], M, g["noises"], float(g["dt_gamma"]), int(g["max_steps"]))
    assert np.array_equal(counter, g["counter"])
    assert np.array_equal(rays, g["rays"])                      # the oracle places segments in ray order, like the fixture
    tot = int(g["counts"].sum())
    assert np.array_equal(x[:tot], g["xyzs"]) and np.array_equal(d[:tot], g["dirs"]) and np.array_equal(l[:tot], g["deltas"])
    ws, dep, img = oracle.composite_rays_train_forward(g["sigmas"], g["rgbs"], g["deltas"], g["rays"], float(g["T_thresh"]))
    assert np.allclose(ws, g["weights_sum"], rtol=1e-5, atol=1e-6) and np.allclose(dep, g["depth"], rtol=1e-5, atol=1e-5)
    assert np.allclose(img, g["image"], rtol=1e-5, atol=1e-6)
    gs, gc = oracle.composite_rays_train_backward(g["grad_weights_sum"], g["grad_depth"], g["grad_image"], g["sigmas"], g["rgbs"], g["deltas"],
                                                  g["rays"], g["weights_sum"], g["depth"], g["image"], float(g["T_thresh"]))
    assert np.allclose(gs, g["grad_sigmas"], rtol=1e-4, atol=1e-5 * np.abs(g["grad_sigmas"]).max())
    assert np.allclose(gc, g["grad_rgbs"], rtol=1e-5, atol=1e-6)


def test_oracle_inference_pair_vs_golden(oracle):
    g = load("march_infer")
    N = g["rays_o"].shape[0]
    ws, dep, img = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, 3), np.float32)
    for it in range(2):
        alive, t = g[f"alive_in{it}"].copy(), g[f"t_in{it}"].copy()
        n_alive, n_step = alive.shape[0], int(g["n_step"])
        x, d, l = oracle.march_rays(n_alive, n_step, alive, t, g["rays_o"], g["rays_d"], float(g["bound"]), g["bitfield"], int(g["C"]), int(g["H"]),
                                    g["nears"], g["fars"], dt_gamma=float(g["dt_gamma"]), max_steps=int(g["max_steps"]))
        assert np.array_equal(x, g[f"xyzs{it}"]) and np.array_equal(d, g[f"dirs{it}"]) and np.array_equal(l, g[f"deltas{it}"])
        oracle.composite_rays(n_alive, n_step, alive, t, g[f"sigmas{it}"], g[f"rgbs{it}"], l, ws, dep, img, float(g["T_thresh"]))
        assert np.array_equal(alive, g[f"alive_out{it}"])
        assert np.allclose(t, g[f"t_out{it}"], rtol=1e-6, atol=1e-6)
        assert np.allclose(ws, g[f"ws{it}"], rtol=1e-5, atol=1e-6) and np.allclose(img, g[f"image{it}"], rtol=1e-5, atol=1e-6)
        assert np.allclose(dep, g[f"depth{it}"], rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("name", ["ffmlp_32_64x2_16", "ffmlp_32_64x3_16"])
def test_oracle_ffmlp_vs_golden(oracle, name):
    """The reference accumulates in fp16 (wmma half accumulators); the oracle in fp32 with fp16 activations: 5e-3 of scale."""
    g = load(name)
    i, o, h, nl = [int(v) for v in g["dims"]]
    y, acts = oracle.ffmlp_forward(g["inputs"], g["weights"], i, o, h, nl, round_activations_fp16=True)
    assert rel(y, g["outputs"]) < 5e-3 and rel(y, g["outputs_inference"]) < 5e-3
    for l in range(nl):
        assert rel(acts[l], g["forward_buffer"][l]) < 5e-3
    gx, gw = oracle.ffmlp_backward(g["grad"], g["inputs"], g["weights"], i, o, h, nl, round_activations_fp16=True)
    # relative Frobenius: an fp16-accumulated pre-activation next to zero flips a ReLU mask (rank-1 outliers)
    fro = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
    assert fro(gx, g["grad_inputs"]) < 3e-2 and fro(gw, g["grad_weights"]) < 3e-2


def test_oracle_sh_vs_golden(oracle):
    g = load("shencoder_deg4")
    assert np.allclose(oracle.sh_encode(g["dirs"], 4), g["outputs"], rtol=1e-6, atol=1e-6)
    g = load("shencoder_deg8")  # bands 4..7: the reference evaluates expanded fp32 polynomials with coefficients up to ~700
    o = oracle.sh_encode(g["dirs"], 8)
    assert np.allclose(o[:, :16], g["outputs"][:, :16], rtol=1e-6, atol=1e-6)
    assert np.abs(o - g["outputs"]).max() <= 1e-5 * max(1.0, np.abs(o).max())


def test_oracle_grid_tv_vs_golden(oracle):
    """R4 pinned without a GPU: oracle.grad_total_variation against the reference extension's output (fp32; its half path is a
    no-op, see tests/golden/make_golden.py::grid_tv)."""
    g = load("grid_tv")
    oracle.set_level_exp2(g["level_exp2"])
    try:
        grad = np.zeros_like(g["embeddings_f32"])
        oracle.grad_total_variation(g["x_f32"], g["embeddings_f32"], grad, g["offsets"], float(g["weight"]), float(g["per_level_scale"]),
                                    int(g["base_resolution"]))
    finally:
        oracle.set_level_exp2(None)
    assert np.abs(g["grad_f32"]).max() > 0
    assert rel(grad, g["grad_f32"]) < 1e-5


# ---------------------------------------------------------------------------------------------- kernels vs golden (GPU)
def _T(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("tag,tol", [("f32", 1e-5), ("f16", 2e-3)])
def test_kernels_grid_vs_golden(tag, tol):
    import torch
    import gridencoder
    import ngp_abi as abi
    g = load("grid_" + tag)
    be = gridencoder._backend
    dt = torch.float32 if tag == "f32" else torch.float16
    x, emb, offs = _T(g["x"]), _T(g["embeddings"]).to(dt), _T(g["offsets"])
    L, B, C = g["outputs"].shape
    S, H = float(np.log2(float(g["per_level_scale"]))), int(g["base_resolution"])
    out = torch.empty(L, B, C, device="cuda", dtype=dt)
    dy = torch.empty(B, L * 3 * C, device="cuda", dtype=dt)
    # default call, no option of any kind: the library's dy_dx is the reference's (gridencoder.cu:143)
    be.grid_encode_forward(x, emb, offs, out, B, 3, C, L, S, H, dy, 0, False, 0)
    ge = torch.zeros_like(emb)
    gi = torch.zeros(B, 3, device="cuda", dtype=dt)
    be.grid_encode_backward(_T(g["grad"]).to(dt), x, emb, offs, ge, B, 3, C, L, S, H, dy, gi, 0, False, 0)
    assert rel(out.float().cpu().numpy(), g["outputs"]) < tol
    assert rel(ge.float().cpu().numpy(), g["grad_embeddings"]) < tol * 4
    assert rel(gi.float().cpu().numpy(), g["grad_inputs"]) < max(tol * 10, 1e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["g0", "g128"])
def test_kernels_march_and_composite_vs_golden(tag):
    import torch
    import raymarching as rm
    g = load("march_train_" + tag)
    N = g["rays_o"].shape[0]
    M = N * 256
    xyzs, dirs, deltas = torch.zeros(M, 3, device="cuda"), torch.zeros(M, 3, device="cuda"), torch.zeros(M, 2, device="cuda")
    rays = torch.empty(N, 3, dtype=torch.int32, device="cuda")
    counter = torch.zeros(2, dtype=torch.int32, device="cuda")
    rm._backend.march_rays_train(_T(g["rays_o"]), _T(g["rays_d"]), _T(g["bitfield"]), float(g["bound"]), float(g["dt_gamma"]), int(g["max_steps"]), N,
                                 int(g["C"]), int(g["H"]), M, _T(g["nears"]), _T(g["fars"]), xyzs, dirs, deltas, rays, counter, _T(g["noises"]))
    tot = int(g["counts"].sum())
    assert np.array_equal(counter.cpu().numpy(), g["counter"]) and np.array_equal(rays.cpu().numpy(), g["rays"])
    assert np.array_equal(xyzs[:tot].cpu().numpy(), g["xyzs"]) and np.array_equal(deltas[:tot].cpu().numpy(), g["deltas"])
    assert np.array_equal(dirs[:tot].cpu().numpy(), g["dirs"])
    sg, cl = _T(g["sigmas"]).requires_grad_(True), _T(g["rgbs"]).requires_grad_(True)
    ws, dep, img = rm.composite_rays_train(sg, cl, _T(g["deltas"]), _T(g["rays"]), float(g["T_thresh"]))
    assert np.allclose(img.detach().cpu().numpy(), g["image"], rtol=1e-5, atol=1e-6)
    assert np.allclose(ws.detach().cpu().numpy(), g["weights_sum"], rtol=1e-5, atol=1e-6)
    torch.autograd.backward([ws, dep, img], [_T(g["grad_weights_sum"]), _T(g["grad_depth"]), _T(g["grad_image"])])
    assert np.allclose(sg.grad.cpu().numpy(), g["grad_sigmas"], rtol=1e-4, atol=1e-5 * np.abs(g["grad_sigmas"]).max())
    assert np.allclose(cl.grad.cpu().numpy(), g["grad_rgbs"], rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ffmlp_32_64x2_16", "ffmlp_32_64x3_16"])
def test_kernels_ffmlp_vs_golden(name):
    import torch
    import ffmlp
    g = load(name)
    i, o, h, nl = [int(v) for v in g["dims"]]
    be = ffmlp._backend
    x, w, gr = _T(g["inputs"]).half(), _T(g["weights"]), _T(g["grad"]).half()
    B = x.shape[0]
    y = torch.empty(B, o, device="cuda", dtype=torch.half)
    be.ffmlp_forward(x, w, B, i, o, h, nl, 0, 6, None, y)
    assert rel(y.float().cpu().numpy(), g["outputs"]) < 5e-3
    gx, gw = torch.empty_like(x), torch.empty_like(w)
    be.ffmlp_backward(gr, x, w, None, B, i, o, h, nl, 0, 6, True, None, gx, gw)
    fro = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
    assert fro(gx.float().cpu().numpy(), g["grad_inputs"]) < 3e-2 and fro(gw.cpu().numpy(), g["grad_weights"]) < 3e-2
