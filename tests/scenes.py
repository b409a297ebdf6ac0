"""Synthetic inputs shared by tests, smoke() and bench.py (numpy only).

Scene: gyroid shell inside the unit cube, sigma(x) = 40 * [g(x) > 0.6] * [|x|_inf < 0.9],
g = sin8x cos8y + sin8y cos8z + sin8z cos8x; colour c(x) = 0.5 + 0.5 sin(6x + (0,2,4)).
"""
import numpy as np


def expand_bits(v):
    v = np.asarray(v, dtype=np.uint32)
    v = (v * np.uint32(0x00010001)) & np.uint32(0xFF0000FF)
    v = (v * np.uint32(0x00000101)) & np.uint32(0x0F00F00F)
    v = (v * np.uint32(0x00000011)) & np.uint32(0xC30C30C3)
    v = (v * np.uint32(0x00000005)) & np.uint32(0x49249249)
    return v


def morton3d(x, y, z):
    return expand_bits(x) | (expand_bits(y) << np.uint32(1)) | (expand_bits(z) << np.uint32(2))


def density(xyz):
    x, y, z = 8 * xyz[..., 0], 8 * xyz[..., 1], 8 * xyz[..., 2]
    g = np.sin(x) * np.cos(y) + np.sin(y) * np.cos(z) + np.sin(z) * np.cos(x)
    inside = np.max(np.abs(xyz), axis=-1) < 0.9
    return (40.0 * (g > 0.6) * inside).astype(np.float32)


def colour(xyz):
    return (0.5 + 0.5 * np.sin(6 * xyz + np.array([0.0, 2.0, 4.0]))).astype(np.float32)


def density_grid(H=128, cascade=1, bound=1.0):
    """[cascade, H^3] float32 densities at cell centres, cells in Morton order (nerf/renderer.py:510-538)."""
    idx = np.arange(H, dtype=np.uint32)
    X, Y, Z = np.meshgrid(idx, idx, idx, indexing="ij")
    m = morton3d(X.ravel(), Y.ravel(), Z.ravel()).astype(np.int64)
    grid = np.zeros((cascade, H ** 3), np.float32)
    for cas in range(cascade):
        b = min(2.0 ** cas, bound)
        centres = (np.stack([X.ravel(), Y.ravel(), Z.ravel()], -1).astype(np.float32) + 0.5) / H * 2 - 1
        grid[cas, m] = density(centres * b)
    return grid


def bitfield_from_grid(grid, thresh=0.01):
    """numpy reference of packbits: LSB-first, strict '>' (raymarching.cu:267-289)."""
    return np.packbits((grid.reshape(-1) > thresh), bitorder="little")


def camera_rays(n_rays, seed=0, radius=1.6, fovy_deg=50.0, W=800, Hh=800):
    """Random pixels of random look-at-origin cameras on a sphere (get_rays semantics, nerf/utils.py:52-136)."""
    rng = np.random.default_rng(seed)
    v = rng.normal(size=3)
    v /= np.linalg.norm(v)
    eye = (radius * v).astype(np.float64)
    fwd = -eye / np.linalg.norm(eye)
    up = np.array([0.0, 1.0, 0.0])
    if abs(fwd @ up) > 0.99:
        up = np.array([1.0, 0.0, 0.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    upv = np.cross(right, fwd)
    focal = Hh / (2 * np.tan(np.radians(fovy_deg) / 2))
    i = rng.integers(0, W, n_rays)
    j = rng.integers(0, Hh, n_rays)
    xs = (i + 0.5 - W / 2) / focal
    ys = (j + 0.5 - Hh / 2) / focal
    d = xs[:, None] * right + (-ys)[:, None] * upv + fwd
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    o = np.broadcast_to(eye, d.shape)
    return np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)


def frame_rays(W, Hh, pose_index=0, n_poses=120, radius=1.6, fovy_deg=50.0, rows=None):
    """All rays of one orbit frame, row-major; rows=(r0, r1) selects a band (tile sharding)."""
    ang = 2 * np.pi * pose_index / n_poses
    eye = np.array([radius * np.cos(ang) * 0.9, radius * 0.3, radius * np.sin(ang) * 0.9])
    fwd = -eye / np.linalg.norm(eye)
    up = np.array([0.0, 1.0, 0.0])
    right = np.cross(fwd, up)
    right /= np.linalg.norm(right)
    upv = np.cross(right, fwd)
    focal = Hh / (2 * np.tan(np.radians(fovy_deg) / 2))
    r0, r1 = (0, Hh) if rows is None else rows
    jj, ii = np.meshgrid(np.arange(r0, r1), np.arange(W), indexing="ij")
    xs = (ii.ravel() + 0.5 - W / 2) / focal
    ys = (jj.ravel() + 0.5 - Hh / 2) / focal
    d = xs[:, None] * right + (-ys)[:, None] * upv + fwd
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    o = np.broadcast_to(eye, d.shape)
    return np.ascontiguousarray(o, np.float32), np.ascontiguousarray(d, np.float32)
