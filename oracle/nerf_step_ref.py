"""CPU restatement of ONE `main_nerf -O --ff` training step -- TEST INFRASTRUCTURE (oracle), never on the product path.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may import this.
Composes the per-op oracles of oracle/cpu_ref.py (C + OpenMP for the hash grid, marcher and compositor; numpy for the
MLPs) in the order of /root/reference/nerf/renderer.py::run_cuda (:257-431, training branch :293-345),
nerf/network_ff.py::forward (:51-74), nerf/utils.py::train_step (:497-590: MSE over rgb, mean) and the optimizer step
(main_nerf.py:194 Adam(betas=(0.9, 0.99), eps=1e-15), GradScaler unscale) -- the same sequence
torch-ngp_b200/nerf/trainer.py launches on the device.  Parity pinned through its parts: each per-op oracle is
checked against the reference's own CUDA extensions (tests/test_vs_reference_ext.py, tests/golden/).
"""
import numpy as np

from . import cpu_ref


def _h(x, dtype=np.float32):
    """fp16 rounding point (autocast boundary)."""
    return np.asarray(x, np.float32).astype(np.float16).astype(dtype)


class RefState:
    """Parameters + Adam moments + loss scale, numpy."""

    def __init__(self, table, sigma_w, color_w, offsets, per_level_scale, base_resolution, lr=1e-2, loss_scale=65536.0):  # GradScaler's init_scale (nerf/utils.py:328)
        self.table, self.sigma_w, self.color_w = table.astype(np.float32), sigma_w.astype(np.float32), color_w.astype(np.float32)
        self.offsets, self.pls, self.H0 = np.asarray(offsets, np.int32), float(per_level_scale), int(base_resolution)
        self.m = [np.zeros_like(p) for p in self.params()]
        self.v = [np.zeros_like(p) for p in self.params()]
        self.t, self.lr0, self.scale = 0, lr, loss_scale
        self._h16 = None

    def params(self):
        return [self.table, self.sigma_w, self.color_w]

    def half(self):
        """fp16-rounded copies of the parameters (what the kernels read under autocast), cached until the next update."""
        if self._h16 is None:
            self._h16 = (_h(self.table), _h(self.sigma_w), _h(self.color_w))
        return self._h16


def network_forward(st, xyzs, dirs, bound, density_scale=1.0):
    """-> sigmas [M], rgbs [M,3] and the saved tensors for the backward (network_ff.py:51-74 under fp16 autocast)."""
    x01 = ((xyzs + np.float32(bound)) / np.float32(2 * bound)).astype(np.float32)
    table16, sigma16, color16 = st.half()
    enc_lbc, _ = cpu_ref.grid_encode_forward(x01, table16, st.offsets, st.pls, st.H0)
    M = xyzs.shape[0]
    enc = _h(np.transpose(enc_lbc, (1, 0, 2)).reshape(M, -1))
    h, _ = cpu_ref.ffmlp_forward(enc, sigma16, 32, 16, 64, 2, dtype=np.float32, round_activations_fp16=True)
    h = _h(h)
    sigmas = (np.float32(density_scale) * np.exp(h[:, 0])).astype(np.float32)
    cin = np.zeros((M, 32), np.float32)
    cin[:, :16] = _h(cpu_ref.sh_encode(dirs, 4))
    cin[:, 16:31] = h[:, 1:]
    c, _ = cpu_ref.ffmlp_forward(cin, color16, 32, 16, 64, 3, dtype=np.float32, round_activations_fp16=True)
    c = _h(c)
    rgbs = _h(1.0 / (1.0 + np.exp(-c[:, :3])))
    return sigmas, rgbs, dict(x01=x01, enc=enc, h=h, cin=cin, c=c)


def forward_backward(st, rays_o, rays_d, gt, bitfield, cascade, grid_size, M, bound=1.0, min_near=0.2, dt_gamma=0.0, max_steps=1024,
                     noises=None, bg_color=1.0, T_thresh=1e-4, density_scale=1.0):
    """-> loss, (g_table [T,C], g_sigma_w, g_color_w) carrying the loss scale, n_samples."""
    N = rays_o.shape[0]
    aabb = np.array([-bound] * 3 + [bound] * 3, np.float32)
    nears, fars = cpu_ref.near_far_from_aabb(rays_o, rays_d, aabb, min_near)
    xyzs, dirs, deltas, rays, counter = cpu_ref.march_rays_train(rays_o, rays_d, bound, bitfield, cascade, grid_size, nears, fars, M, noises,
                                                                 dt_gamma, max_steps)
    sigmas, rgbs, sv = network_forward(st, xyzs, dirs, bound, density_scale)
    ws, depth, image = cpu_ref.composite_rays_train_forward(sigmas, rgbs, deltas, rays, T_thresh)
    pred = image + (1 - ws)[:, None] * np.float32(bg_color)
    diff = (pred - gt).astype(np.float32)
    loss = float(np.mean(diff.astype(np.float64) ** 2))
    g_image = (np.float32(st.scale) * 2.0 / (3.0 * N) * diff).astype(np.float32)
    g_ws = -(g_image.sum(-1) * np.float32(bg_color)).astype(np.float32)
    g_sig, g_rgb = cpu_ref.composite_rays_train_backward(g_ws, np.zeros(N, np.float32), g_image, sigmas, rgbs, deltas, rays, ws, depth, image,
                                                         T_thresh)
    g_c = np.zeros((M, 16), np.float32)
    g_c[:, :3] = _h(g_rgb * rgbs * (1 - rgbs))
    g_cin, g_col = cpu_ref.ffmlp_backward(g_c, sv["cin"], st.half()[2], 32, 16, 64, 3, dtype=np.float32, round_activations_fp16=True)
    g_h = np.zeros((M, 16), np.float32)
    g_h[:, 0] = g_sig * np.float32(density_scale) * np.exp(np.clip(sv["h"][:, 0], -15, 15))
    g_h[:, 1:] = _h(g_cin)[:, 16:31]
    g_h = _h(g_h)
    g_enc, g_sw = cpu_ref.ffmlp_backward(g_h, sv["enc"], st.half()[1], 32, 16, 64, 2, dtype=np.float32, round_activations_fp16=True)
    g_enc = _h(g_enc)
    L = st.offsets.shape[0] - 1
    g_lbc = np.ascontiguousarray(np.transpose(g_enc.reshape(M, L, -1), (1, 0, 2)))
    g_table, _ = cpu_ref.grid_encode_backward(g_lbc, sv["x01"], st.half()[0], st.offsets, st.pls, st.H0)
    return loss, (g_table, g_sw.astype(np.float32), g_col.astype(np.float32)), int(counter[0])


def adam_step(st, grads, grad_div=1.0, beta1=0.9, beta2=0.99, eps=1e-15, decay_steps=30000):
    """torch.optim.Adam + LambdaLR(0.1 ** min(iter / 30000, 1)) + GradScaler.unscale_ / skip on non-finite gradients."""
    it = getattr(st, "it", 0)  # iterations so far: the reference's LambdaLR steps every iteration, skipped or not (nerf/utils.py:974-981)
    st.it = it + 1
    if not all(np.isfinite(g).all() for g in grads):
        st.scale *= 0.5
        return False
    lr = st.lr0 * 0.1 ** min(it / decay_steps, 1.0)
    st.t += 1
    st._h16 = None
    bc1, bc2 = 1 - beta1 ** st.t, 1 - beta2 ** st.t
    for p, m, v, g in zip(st.params(), st.m, st.v, grads):
        g = g.reshape(p.shape).astype(np.float32) / np.float32(st.scale * grad_div)
        m *= beta1
        m += (1 - beta1) * g
        v *= beta2
        v += (1 - beta2) * g * g
        p -= (lr / bc1) * m / (np.sqrt(v) / np.sqrt(bc2) + eps)
    return True


def train_step(st, rays_o, rays_d, gt, bitfield, cascade, grid_size, M, **kw):
    loss, grads, n = forward_backward(st, rays_o, rays_d, gt, bitfield, cascade, grid_size, M, **kw)
    adam_step(st, grads)
    return loss, n


def render(st, rays_o, rays_d, bitfield, cascade, grid_size, bound=1.0, min_near=0.2, dt_gamma=0.0, max_steps=1024, bg_color=1.0, T_thresh=1e-4,
           density_scale=1.0):
    """Inference loop of nerf/renderer.py:346-431 on the per-op oracles -> image [N,3], depth [N]."""
    N = rays_o.shape[0]
    aabb = np.array([-bound] * 3 + [bound] * 3, np.float32)
    nears, fars = cpu_ref.near_far_from_aabb(rays_o, rays_d, aabb, min_near)
    ws, depth, image = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, 3), np.float32)
    alive = np.arange(N, dtype=np.int32)
    rays_t = nears.copy()
    step = 0
    while step < max_steps and alive.size > 0:
        n_alive = alive.size
        n_step = max(min(N // n_alive, 8), 1)
        xyzs, dirs, deltas = cpu_ref.march_rays(n_alive, n_step, alive, rays_t, rays_o, rays_d, bound, bitfield, cascade, grid_size, nears, fars,
                                                dt_gamma=dt_gamma, max_steps=max_steps)
        sigmas, rgbs, _ = network_forward(st, xyzs, dirs, bound, density_scale)
        cpu_ref.composite_rays(n_alive, n_step, alive, rays_t, sigmas, rgbs, deltas, ws, depth, image, T_thresh)
        alive = alive[alive >= 0]
        step += n_step
    image = image + (1 - ws)[:, None] * np.float32(bg_color)
    depth = np.clip(depth - nears, 0, None) / (fars - nears)
    return image, depth


# ---- SDF step (main_sdf.py --fp16 --ff): hash grid -> FFMLP 32 -> 64 x 3 -> 1, mape loss ------------------------------------
class SDFRefState:
    def __init__(self, table, mlp_w, offsets, per_level_scale, base_resolution, lr=1e-4, loss_scale=65536.0, weight_decay=1e-6):
        self.table, self.mlp_w = table.astype(np.float32), mlp_w.astype(np.float32)
        self.offsets, self.pls, self.H0 = np.asarray(offsets, np.int32), float(per_level_scale), int(base_resolution)
        self.m, self.v = [np.zeros_like(self.table), np.zeros_like(self.mlp_w)], [np.zeros_like(self.table), np.zeros_like(self.mlp_w)]
        self.t, self.lr, self.scale, self.wd = 0, lr, loss_scale, weight_decay


def sdf_forward_backward(st, points, sdfs, hid=64, nl=3):
    """-> loss, (g_table, g_mlp) carrying the loss scale (sdf/utils.py:217-225 + loss.py:7-15, fp16 autocast)."""
    B = points.shape[0]
    x01 = ((points + np.float32(1)) / np.float32(2)).astype(np.float32)
    t16, w16 = _h(st.table), _h(st.mlp_w)
    enc_lbc, _ = cpu_ref.grid_encode_forward(x01, t16, st.offsets, st.pls, st.H0)
    enc = _h(np.transpose(enc_lbc, (1, 0, 2)).reshape(B, -1))
    h, _ = cpu_ref.ffmlp_forward(enc, w16, 32, 16, hid, nl, dtype=np.float32, round_activations_fp16=True)
    pred = _h(h)[:, 0]
    w = 1.0 / (np.abs(sdfs) + np.float32(1e-2))
    d = pred - sdfs
    loss = float(np.mean(np.abs(d) * w))
    g_h = np.zeros((B, 16), np.float32)
    g_h[:, 0] = _h(np.float32(st.scale) * np.sign(d) * w / np.float32(B))
    g_enc, g_w = cpu_ref.ffmlp_backward(g_h, enc, w16, 32, 16, hid, nl, dtype=np.float32, round_activations_fp16=True)
    L = st.offsets.shape[0] - 1
    g_lbc = np.ascontiguousarray(np.transpose(_h(g_enc).reshape(B, L, -1), (1, 0, 2)))
    g_table, _ = cpu_ref.grid_encode_backward(g_lbc, x01, t16, st.offsets, st.pls, st.H0)
    return loss, (g_table, g_w.astype(np.float32))


def sdf_adam_step(st, grads, grad_div=1.0, beta1=0.9, beta2=0.99, eps=1e-15):
    if not all(np.isfinite(g).all() for g in grads):
        st.scale *= 0.5
        return False
    st.t += 1
    bc1, bc2 = 1 - beta1 ** st.t, 1 - beta2 ** st.t
    for k, (p, g) in enumerate(zip((st.table, st.mlp_w), grads)):
        g = g.reshape(p.shape).astype(np.float32) / np.float32(st.scale * grad_div) + (np.float32(st.wd) * p if k == 1 else 0)
        st.m[k] = beta1 * st.m[k] + (1 - beta1) * g
        st.v[k] = beta2 * st.v[k] + (1 - beta2) * g * g
        p -= (st.lr / bc1) * st.m[k] / (np.sqrt(st.v[k]) / np.sqrt(bc2) + eps)
    return True
