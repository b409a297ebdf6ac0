"""TEST INFRASTRUCTURE ONLY -- numpy/ctypes front-end of the CPU oracle.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Nothing under torch-ngp_b200/ does (the product path has no CPU fallback).

* hash grid + ray marching: thin ctypes wrappers over oracle/ngp_oracle.c (built by
  ``build()`` below with gcc; each C function cites the reference lines it restates).
* FFMLP: numpy restatement of the bias-free MLP twin of /root/reference/testing/test_ffmlp.py:11-43
  with the weight layout of /root/reference/ffmlp/ffmlp.py:117-122.
* SH direction encoding: numpy restatement of /root/reference/shencoder/src/shencoder.cu:27-356
  (degree <= 4 is what the NeRF networks use, network_ff.py:44).

Parity pinning: see the header of ngp_oracle.c.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "ngp_oracle.c")
_LIB_DIR = os.path.join(_HERE, "_build")
_LIB = os.path.join(_LIB_DIR, "libngp_oracle.so")
_lib = None

_u32, _f32, _i32 = ctypes.c_uint32, ctypes.c_float, ctypes.c_int
_p = ctypes.c_void_p


def build(force=False):
    """gcc the C restatement (no-op when up to date)."""
    if not force and os.path.exists(_LIB) and os.path.getmtime(_LIB) >= os.path.getmtime(_SRC):
        return _LIB
    os.makedirs(_LIB_DIR, exist_ok=True)
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC", _SRC, "-o", _LIB, "-lm"]
    subprocess.check_call(cmd)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def set_pos_deriv_quirk(on):
    """1 (the default state): the reference's `pos_deriv[D] = {1.0f}` (gridencoder.cu:143, see ngp_oracle.c); 0: the exact
    derivative, for checking the kernels' exact_input_grad opt-in.  Callers that switch it off must switch it back on."""
    lib().orc_set_pos_deriv_quirk(_i32(int(on)))


_exp2_keepalive = None


def set_level_exp2(values):
    """Per-level exp2f(level * S) as evaluated by the GPU (float32 array), or None for libm (see ngp_oracle.c)."""
    global _exp2_keepalive
    if values is None:
        _exp2_keepalive = None
        lib().orc_set_level_exp2(None, _u32(0))
    else:
        _exp2_keepalive = np.ascontiguousarray(values, dtype=np.float32)
        lib().orc_set_level_exp2(_exp2_keepalive.ctypes.data_as(_p), _u32(_exp2_keepalive.size))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_p)


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


# ------------------------------------------------------------------------------------------------
# hash grid
# ------------------------------------------------------------------------------------------------
def make_offsets(input_dim=3, num_levels=16, level_dim=2, per_level_scale=2, base_resolution=16,
                 log2_hashmap_size=19, desired_resolution=None, align_corners=False):
    """Level table sizing, /root/reference/gridencoder/grid.py:758-786 (float64 host arithmetic)."""
    if desired_resolution is not None:
        per_level_scale = np.exp2(np.log2(desired_resolution / base_resolution) / (num_levels - 1))
    offsets, offset, max_params = [], 0, 2 ** log2_hashmap_size
    for i in range(num_levels):
        resolution = int(np.ceil(base_resolution * per_level_scale ** i))
        n = min(max_params, (resolution if align_corners else resolution + 1) ** input_dim)
        n = int(np.ceil(n / 8) * 8)
        offsets.append(offset)
        offset += n
    offsets.append(offset)
    return np.array(offsets, dtype=np.int32), float(per_level_scale)


def grid_encode_forward(inputs, embeddings, offsets, per_level_scale, base_resolution, calc_grad_inputs=False,
                        gridtype=0, align_corners=False, interpolation=0):
    """-> (outputs [L,B,C] level-major like the reference kernel, dy_dx [B,L*D*C] or None)."""
    dt = np.float64 if embeddings.dtype == np.float64 else np.float32
    sfx = "f64" if dt == np.float64 else "f32"
    inputs = _c(inputs, np.float32)
    emb = _c(embeddings, dt)
    offsets = _c(offsets, np.int32)
    B, D = inputs.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    S = np.float32(np.log2(per_level_scale))
    out = np.empty((L, B, C), dtype=dt)
    dy_dx = np.empty((B, L * D * C), dtype=dt) if calc_grad_inputs else None
    getattr(lib(), f"orc_grid_encode_forward_{sfx}")(
        _ptr(inputs), _ptr(emb), _ptr(offsets), _ptr(out), _u32(B), _u32(D), _u32(C), _u32(L), _f32(S),
        _u32(base_resolution), _ptr(dy_dx), _u32(gridtype), _i32(int(align_corners)), _u32(interpolation))
    return out, dy_dx


def grid_encode_backward(grad, inputs, embeddings, offsets, per_level_scale, base_resolution, dy_dx=None,
                         gridtype=0, align_corners=False, interpolation=0):
    """grad: [L,B,C].  -> (grad_embeddings [T,C], grad_inputs [B,D] or None)."""
    dt = np.float64 if embeddings.dtype == np.float64 else np.float32
    sfx = "f64" if dt == np.float64 else "f32"
    inputs = _c(inputs, np.float32)
    emb = _c(embeddings, dt)
    grad = _c(grad, dt)
    offsets = _c(offsets, np.int32)
    B, D = inputs.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    S = np.float32(np.log2(per_level_scale))
    g_emb = np.zeros_like(emb)
    g_in = np.zeros((B, D), dtype=dt) if dy_dx is not None else None
    dy = _c(dy_dx, dt) if dy_dx is not None else None
    getattr(lib(), f"orc_grid_encode_backward_{sfx}")(
        _ptr(grad), _ptr(inputs), _ptr(emb), _ptr(offsets), _ptr(g_emb), _u32(B), _u32(D), _u32(C), _u32(L),
        _f32(S), _u32(base_resolution), _ptr(dy), _ptr(g_in), _u32(gridtype), _i32(int(align_corners)),
        _u32(interpolation))
    return g_emb, g_in


def grad_total_variation(inputs, embeddings, grad, offsets, weight, per_level_scale, base_resolution,
                         gridtype=0, align_corners=False):
    """In-place on ``grad`` ([T,C], same dtype as embeddings)."""
    dt = np.float64 if embeddings.dtype == np.float64 else np.float32
    sfx = "f64" if dt == np.float64 else "f32"
    inputs = _c(inputs, dt)
    emb = _c(embeddings, dt)
    assert grad.dtype == dt and grad.flags.c_contiguous
    offsets = _c(offsets, np.int32)
    B, D = inputs.shape
    L, C = offsets.shape[0] - 1, emb.shape[1]
    S = np.float32(np.log2(per_level_scale))
    getattr(lib(), f"orc_grad_total_variation_{sfx}")(
        _ptr(inputs), _ptr(emb), _ptr(grad), _ptr(offsets), _f32(weight), _u32(B), _u32(D), _u32(C), _u32(L),
        _f32(S), _u32(base_resolution), _u32(gridtype), _i32(int(align_corners)))
    return grad


# ------------------------------------------------------------------------------------------------
# ray marching
# ------------------------------------------------------------------------------------------------
def near_far_from_aabb(rays_o, rays_d, aabb, min_near=0.2):
    rays_o, rays_d, aabb = _c(rays_o, np.float32).reshape(-1, 3), _c(rays_d, np.float32).reshape(-1, 3), _c(aabb, np.float32)
    N = rays_o.shape[0]
    nears, fars = np.empty(N, np.float32), np.empty(N, np.float32)
    lib().orc_near_far_from_aabb(_ptr(rays_o), _ptr(rays_d), _ptr(aabb), _u32(N), _f32(min_near), _ptr(nears), _ptr(fars))
    return nears, fars


def sph_from_ray(rays_o, rays_d, radius):
    rays_o, rays_d = _c(rays_o, np.float32).reshape(-1, 3), _c(rays_d, np.float32).reshape(-1, 3)
    N = rays_o.shape[0]
    coords = np.empty((N, 2), np.float32)
    lib().orc_sph_from_ray(_ptr(rays_o), _ptr(rays_d), _f32(radius), _u32(N), _ptr(coords))
    return coords


def morton3D(coords):
    coords = _c(coords, np.int32)
    N = coords.shape[0]
    out = np.empty(N, np.int32)
    lib().orc_morton3D(_ptr(coords), _u32(N), _ptr(out))
    return out


def morton3D_invert(indices):
    indices = _c(indices, np.int32)
    N = indices.shape[0]
    out = np.empty((N, 3), np.int32)
    lib().orc_morton3D_invert(_ptr(indices), _u32(N), _ptr(out))
    return out


def mark_untrained_grid(poses, intrinsic, bound, cascade, H):
    """count[C, H^3] (Morton order) of the cameras that see each cell -- numpy restatement of
    NeRFRenderer.mark_untrained_grid (nerf/renderer.py:433-496) without its blocking loops, float32 with one rounding per
    torch op; the camera-space product is summed ((c0 R0j + c1 R1j) + c2 R2j).  Cells with count 0 are the ones the
    reference sets to -1 (:493)."""
    f = np.float32
    poses = _c(poses, np.float32).reshape(-1, 4, 4)
    fx, fy, cx, cy = [float(v) for v in intrinsic]
    r = np.arange(H, dtype=np.int32)
    xx, yy, zz = np.meshgrid(r, r, r, indexing="ij")
    coords = np.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], -1).astype(np.int32)  # :460
    indices = morton3D(coords).astype(np.int64)                                                # :461
    world = (f(2) * coords.astype(np.float32) / f(H - 1) - f(1)).astype(np.float32)            # :462
    count = np.zeros((cascade, H ** 3), np.int32)
    kx, ky = f(cx / fx), f(cy / fy)
    for cas in range(cascade):
        b = min(2 ** cas, bound)                                                               # :466
        hgs = b / H                                                                            # :467
        cw = (world * f(b - hgs)).astype(np.float32)                                           # :469
        margin = f(hgs * 2)
        acc = np.zeros(H ** 3, np.int32)
        for P in poses:                                                                        # :473-486, one camera at a time
            d = (cw - P[:3, 3][None, :]).astype(np.float32)
            cam = [((d[:, 0] * P[0, j]).astype(np.float32) + (d[:, 1] * P[1, j]).astype(np.float32)).astype(np.float32) for j in range(3)]
            cam = [(cam[j] + (d[:, 2] * P[2, j]).astype(np.float32)).astype(np.float32) for j in range(3)]
            m = (cam[2] > 0) & (np.abs(cam[0]) < (kx * cam[2]).astype(np.float32) + margin) & (np.abs(cam[1]) < (ky * cam[2]).astype(np.float32) + margin)
            acc += m.astype(np.int32)
        count[cas, indices] = acc                                                              # :489
    return count


def packbits(grid, thresh):
    grid = _c(grid, np.float32)
    N = grid.size // 8
    out = np.empty(N, np.uint8)
    lib().orc_packbits(_ptr(grid), _u32(N), _f32(thresh), _ptr(out))
    return out


def density_grid_update(density_grid, tmp_grid, decay=0.95, density_thresh=10.0):
    """The tail of NeRFRenderer.update_extra_state, /root/reference/nerf/renderer.py:582-590, on given grids:
        valid = (grid >= 0) & (tmp >= 0)                         :582
        grid[valid] = maximum(grid[valid] * decay, tmp[valid])   :583   (fp32, like the torch tensors)
        mean_density = mean(clamp(grid, min=0))                  :584   (-1 cells count as 0)
        thresh = min(mean_density, density_thresh)               :589
        bitfield = packbits(grid, thresh)                        :590   (strict `>`, raymarching.cu:292-313)
    -> (new grid [C, H^3] f32, mean_density (python float of the fp32 mean), bitfield uint8 [C * H^3 / 8]).
    The mean is accumulated in float64 and rounded to fp32 once: the tie-breaker for fp32 reductions in any order."""
    grid = np.array(density_grid, np.float32, copy=True)
    tmp = np.asarray(tmp_grid, np.float32)
    valid = (grid >= 0) & (tmp >= 0)
    grid[valid] = np.maximum(grid[valid] * np.float32(decay), tmp[valid])
    mean = float(np.float32(np.clip(grid, 0, None).astype(np.float64).mean()))
    thresh = min(mean, float(density_thresh))
    return grid, mean, packbits(grid, thresh)


def march_rays_train(rays_o, rays_d, bound, density_bitfield, C, H, nears, fars, M, noises=None, dt_gamma=0.0,
                     max_steps=1024, counter=None):
    """Raw kernel semantics (no Python-side trimming).  -> xyzs[M,3], dirs[M,3], deltas[M,2], rays[N,3], counter[2]."""
    rays_o, rays_d = _c(rays_o, np.float32).reshape(-1, 3), _c(rays_d, np.float32).reshape(-1, 3)
    grid = _c(density_bitfield, np.uint8)
    nears, fars = _c(nears, np.float32), _c(fars, np.float32)
    N = rays_o.shape[0]
    noises = np.zeros(N, np.float32) if noises is None else _c(noises, np.float32)
    xyzs, dirs, deltas = np.zeros((M, 3), np.float32), np.zeros((M, 3), np.float32), np.zeros((M, 2), np.float32)
    rays = np.zeros((N, 3), np.int32)
    counter = np.zeros(2, np.int32) if counter is None else counter
    lib().orc_march_rays_train(_ptr(rays_o), _ptr(rays_d), _ptr(grid), _f32(bound), _f32(dt_gamma), _u32(max_steps),
                               _u32(N), _u32(C), _u32(H), _u32(M), _ptr(nears), _ptr(fars), _ptr(xyzs), _ptr(dirs),
                               _ptr(deltas), _ptr(rays), _ptr(counter), _ptr(noises))
    return xyzs, dirs, deltas, rays, counter


def composite_rays_train_forward(sigmas, rgbs, deltas, rays, T_thresh=1e-4):
    sigmas, rgbs, deltas, rays = _c(sigmas, np.float32), _c(rgbs, np.float32), _c(deltas, np.float32), _c(rays, np.int32)
    M, N = sigmas.shape[0], rays.shape[0]
    ws, depth, image = np.zeros(N, np.float32), np.zeros(N, np.float32), np.zeros((N, 3), np.float32)
    lib().orc_composite_rays_train_forward(_ptr(sigmas), _ptr(rgbs), _ptr(deltas), _ptr(rays), _u32(M), _u32(N),
                                           _f32(T_thresh), _ptr(ws), _ptr(depth), _ptr(image))
    return ws, depth, image


def composite_rays_train_backward(grad_ws, grad_depth, grad_image, sigmas, rgbs, deltas, rays, ws, depth, image,
                                  T_thresh=1e-4):
    a = [_c(v, np.float32) for v in (grad_ws, grad_depth, grad_image, sigmas, rgbs, deltas)]
    rays = _c(rays, np.int32)
    b = [_c(v, np.float32) for v in (ws, depth, image)]
    M, N = a[3].shape[0], rays.shape[0]
    g_sig, g_rgb = np.zeros(M, np.float32), np.zeros((M, 3), np.float32)
    lib().orc_composite_rays_train_backward(*[_ptr(v) for v in a], _ptr(rays), *[_ptr(v) for v in b], _u32(M), _u32(N),
                                            _f32(T_thresh), _ptr(g_sig), _ptr(g_rgb))
    return g_sig, g_rgb


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, density_bitfield, C, H, nears, fars,
               M=None, noises=None, dt_gamma=0.0, max_steps=1024):
    rays_alive, rays_t = _c(rays_alive, np.int32), _c(rays_t, np.float32)
    rays_o, rays_d = _c(rays_o, np.float32).reshape(-1, 3), _c(rays_d, np.float32).reshape(-1, 3)
    grid = _c(density_bitfield, np.uint8)
    nears, fars = _c(nears, np.float32), _c(fars, np.float32)
    M = n_alive * n_step if M is None else M
    noises = np.zeros(n_alive, np.float32) if noises is None else _c(noises, np.float32)
    xyzs, dirs, deltas = np.zeros((M, 3), np.float32), np.zeros((M, 3), np.float32), np.zeros((M, 2), np.float32)
    lib().orc_march_rays(_u32(n_alive), _u32(n_step), _ptr(rays_alive), _ptr(rays_t), _ptr(rays_o), _ptr(rays_d),
                         _f32(bound), _f32(dt_gamma), _u32(max_steps), _u32(C), _u32(H), _ptr(grid), _ptr(nears),
                         _ptr(fars), _ptr(xyzs), _ptr(dirs), _ptr(deltas), _ptr(noises))
    return xyzs, dirs, deltas


def composite_rays(n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image, T_thresh=1e-2):
    """In place on rays_alive / rays_t / weights_sum / depth / image (all must be C-contiguous numpy arrays)."""
    for a, dt in ((rays_alive, np.int32), (rays_t, np.float32), (weights_sum, np.float32), (depth, np.float32),
                  (image, np.float32)):
        assert a.dtype == dt and a.flags.c_contiguous
    sigmas, rgbs, deltas = _c(sigmas, np.float32), _c(rgbs, np.float32), _c(deltas, np.float32)
    lib().orc_composite_rays(_u32(n_alive), _u32(n_step), _f32(T_thresh), _ptr(rays_alive), _ptr(rays_t), _ptr(sigmas),
                             _ptr(rgbs), _ptr(deltas), _ptr(weights_sum), _ptr(depth), _ptr(image))


# ------------------------------------------------------------------------------------------------
# FFMLP (numpy): ffmlp.py:117-122 weight layout, test_ffmlp.py:38-43 forward
# ------------------------------------------------------------------------------------------------
def ffmlp_split_weights(weights, input_dim, output_dim_padded, hidden_dim, num_layers):
    """flat -> [W0 [hidden,in], W1..W_{n-1} [hidden,hidden], W_last [out_pad,hidden]] (row-major [out,in])."""
    ws, o = [], 0
    ws.append(weights[o:o + hidden_dim * input_dim].reshape(hidden_dim, input_dim)); o += hidden_dim * input_dim
    for _ in range(num_layers - 1):
        ws.append(weights[o:o + hidden_dim * hidden_dim].reshape(hidden_dim, hidden_dim)); o += hidden_dim * hidden_dim
    ws.append(weights[o:o + output_dim_padded * hidden_dim].reshape(output_dim_padded, hidden_dim))
    return ws


_K_ACT = 10.0  # ffmlp/src/utils.h:41


def ffmlp_activation(act, x):
    """hidden activation by the reference's code (ffmlp.cu:22-32, utils.h:425-470)"""
    if act == 0: return np.maximum(x, 0)
    if act == 1: return np.exp(x)
    if act == 2: return np.sin(x)
    if act == 3: return 1.0 / (1.0 + np.exp(-x))
    if act == 4: return 0.5 * (x * _K_ACT + np.sqrt((x * _K_ACT) ** 2 + 4)) / _K_ACT
    if act == 5: return np.log(np.exp(x * _K_ACT) + 1.0) / _K_ACT
    return x


def ffmlp_activation_backward(act, y):
    """d act / d pre-activation from the stored post-activation y (utils.h:538-580)"""
    if act == 0: return (y > 0).astype(y.dtype)
    if act == 1: return y
    if act == 3: return y * (1.0 - y)
    if act == 4: return (y * _K_ACT) ** 2 / ((y * _K_ACT) ** 2 + 1.0)
    if act == 5: return 1.0 - np.exp(-y * _K_ACT)
    if act == 6: return np.ones_like(y)
    raise ValueError("Sine has no backward from post-activations (utils.h:552-556)")


def ffmlp_forward(inputs, weights, input_dim, output_dim_padded, hidden_dim, num_layers, dtype=np.float64,
                  round_activations_fp16=False, activation=0):
    """-> (outputs [B,out_pad], [post-activation of every hidden layer]).  Hidden activation by code (default ReLU), linear output."""
    x = np.asarray(inputs, dtype=dtype)
    acts = []
    ws = ffmlp_split_weights(np.asarray(weights, dtype=dtype), input_dim, output_dim_padded, hidden_dim, num_layers)
    for i, w in enumerate(ws):
        x = x @ w.T
        if i != num_layers:
            x = ffmlp_activation(activation, x)
            if round_activations_fp16:
                x = x.astype(np.float16).astype(dtype)
            acts.append(x)
    return x, acts


def ffmlp_backward(grad, inputs, weights, input_dim, output_dim_padded, hidden_dim, num_layers, dtype=np.float64,
                   round_activations_fp16=False, activation=0):
    """-> (grad_inputs [B,in], grad_weights flat), matching ffmlp.cu:749-895's results.
    round_activations_fp16: the saved activations (and so the ReLU masks, ffmlp.cu:105-108) are the fp16-rounded ones,
    as both the reference and the tcgen05 kernels store them -- a pre-activation within rounding distance of zero
    otherwise flips a mask and changes that unit's whole gradient contribution."""
    x = np.asarray(inputs, dtype=dtype)
    ws = ffmlp_split_weights(np.asarray(weights, dtype=dtype), input_dim, output_dim_padded, hidden_dim, num_layers)
    _, acts = ffmlp_forward(x, weights, input_dim, output_dim_padded, hidden_dim, num_layers, dtype, round_activations_fp16, activation)
    g = np.asarray(grad, dtype=dtype)
    gws = [None] * len(ws)
    for i in range(num_layers, -1, -1):
        a_prev = acts[i - 1] if i > 0 else x
        gws[i] = g.T @ a_prev
        g = g @ ws[i]
        if i > 0:
            g = g * ffmlp_activation_backward(activation, acts[i - 1])
    return g, np.concatenate([w.reshape(-1) for w in gws])


# ------------------------------------------------------------------------------------------------
# SH direction encoding, degree <= 8 (shencoder.cu:27-120): bands l < 4 in the reference's polynomial form, bands 4..7
# from the closed form the reference's table was generated from (its own comments give each entry in closed form,
# e.g. :72 `3*sqrt(35)*xy*(x2 - y2)/(4*sqrt(pi))`): Y_l^{+-m} = N_l^m Q_l^m(z) {Re|Im}(x+iy)^m, Q_l^m = (-1)^m d^m P_l/dz^m
# ------------------------------------------------------------------------------------------------
def sh_encode(dirs, degree=4):
    d = np.asarray(dirs, dtype=np.float64)
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    xy, xz, yz, x2, y2, z2 = x * y, x * z, y * z, x * x, y * y, z * z
    out = np.zeros((d.shape[0], degree * degree), np.float64)
    out[:, 0] = 0.28209479177387814
    if degree > 1:
        out[:, 1] = -0.48860251190291987 * y
        out[:, 2] = 0.48860251190291987 * z
        out[:, 3] = -0.48860251190291987 * x
    if degree > 2:
        out[:, 4] = 1.0925484305920792 * xy
        out[:, 5] = -1.0925484305920792 * yz
        out[:, 6] = 0.94617469575755997 * z2 - 0.31539156525251999
        out[:, 7] = -1.0925484305920792 * xz
        out[:, 8] = 0.54627421529603959 * x2 - 0.54627421529603959 * y2
    if degree > 3:
        out[:, 9] = 0.59004358992664352 * y * (-3.0 * x2 + y2)
        out[:, 10] = 2.8906114426405538 * xy * z
        out[:, 11] = 0.45704579946446572 * y * (1.0 - 5.0 * z2)
        out[:, 12] = 0.3731763325901154 * z * (5.0 * z2 - 3.0)
        out[:, 13] = 0.45704579946446572 * x * (1.0 - 5.0 * z2)
        out[:, 14] = 1.4453057213202769 * z * (x2 - y2)
        out[:, 15] = 0.59004358992664352 * x * (-x2 + 3.0 * y2)
    if degree > 4:
        from math import factorial, pi, sqrt
        w = x + 1j * y
        for m in range(degree):
            wm = w ** m
            qmm = (-1.0) ** m * float(np.prod(np.arange(2 * m - 1, 0, -2))) if m else 1.0
            q2, q1 = None, np.full_like(z, qmm)
            for l in range(m, degree):
                if l == m:
                    q = q1
                elif l == m + 1:
                    q = (2 * m + 1) * z * q1
                else:
                    q = ((2 * l - 1) * z * q1 - (l + m - 1) * q2) / (l - m)
                if l > m:
                    q2, q1 = q1, q
                if l >= 4:
                    n = sqrt((2 * l + 1) / (4 * pi) * factorial(l - m) / factorial(l + m)) * (sqrt(2) if m else 1.0)
                    out[:, l * l + l + m] = n * q * wm.real
                    if m:
                        out[:, l * l + l - m] = n * q * wm.imag
    return out
