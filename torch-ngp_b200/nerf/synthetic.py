"""Synthetic scene + ray pools for tests and bench.py (there are no datasets on the box).

The scene is the gyroid shell of tests/scenes.py (sigma = 40 inside the shell, colour = 0.5 + 0.5 sin(6x + phase)),
restated in torch so that ground-truth pixels can be rendered on the device with the raymarching ops themselves:
march_rays_train over the analytic occupancy grid, analytic sigma / colour at the samples, composite_rays_train.
Cameras follow get_rays (nerf/utils.py:52-136): look-at-origin poses on a sphere, pinhole, random pixels.
"""
import math

import numpy as np
import torch

import raymarching as _raymarching


def density(xyz):
    x, y, z = 8 * xyz[..., 0], 8 * xyz[..., 1], 8 * xyz[..., 2]
    g = torch.sin(x) * torch.cos(y) + torch.sin(y) * torch.cos(z) + torch.sin(z) * torch.cos(x)
    inside = xyz.abs().amax(-1) < 0.9
    return 40.0 * ((g > 0.6) & inside).float()


def colour(xyz):
    phase = torch.tensor([0.0, 2.0, 4.0], device=xyz.device)
    return 0.5 + 0.5 * torch.sin(6 * xyz + phase)


def random_rays(n_views, rays_per_view, device, seed=0, radius=1.6, fovy_deg=50.0, W=800, H=800):
    """[n_views * rays_per_view, 3] origins and unit directions: random pixels of random look-at-origin cameras."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    v = torch.randn(n_views, 3, generator=g, dtype=torch.float64)
    v = v / v.norm(dim=-1, keepdim=True)
    eye = radius * v
    fwd = -v
    up = torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64).expand(n_views, 3).clone()
    near_pole = (fwd * up).sum(-1).abs() > 0.99
    up[near_pole] = torch.tensor([1.0, 0.0, 0.0], dtype=torch.float64)
    right = torch.linalg.cross(fwd, up)
    right = right / right.norm(dim=-1, keepdim=True)
    upv = torch.linalg.cross(right, fwd)
    focal = H / (2 * math.tan(math.radians(fovy_deg) / 2))
    i = torch.randint(0, W, (n_views, rays_per_view), generator=g)
    j = torch.randint(0, H, (n_views, rays_per_view), generator=g)
    xs = ((i + 0.5 - W / 2) / focal).double()
    ys = ((j + 0.5 - H / 2) / focal).double()
    d = xs[..., None] * right[:, None] + (-ys)[..., None] * upv[:, None] + fwd[:, None]
    d = d / d.norm(dim=-1, keepdim=True)
    o = eye[:, None].expand_as(d)
    return o.reshape(-1, 3).float().contiguous().to(device), d.reshape(-1, 3).float().contiguous().to(device)


def frame_rays(W, H, device, pose_index=0, n_poses=120, radius=1.6, fovy_deg=50.0, rows=None):
    """All rays of one orbit frame, row-major; rows=(r0, r1) selects a band, rows=<1-D index tensor / array> the listed
    rows (tile sharding across ranks: nerf/parallel.py row_band / row_stripes)."""
    ang = 2 * math.pi * pose_index / n_poses
    eye = np.array([radius * math.cos(ang) * 0.9, radius * 0.3, radius * math.sin(ang) * 0.9])
    fwd = -eye / np.linalg.norm(eye)
    right = np.cross(fwd, np.array([0.0, 1.0, 0.0]))
    right /= np.linalg.norm(right)
    upv = np.cross(right, fwd)
    focal = H / (2 * math.tan(math.radians(fovy_deg) / 2))
    if rows is None:
        row_idx = np.arange(0, H)
    elif isinstance(rows, tuple):
        row_idx = np.arange(rows[0], rows[1])
    else:
        row_idx = np.asarray(rows, dtype=np.int64)
    jj, ii = np.meshgrid(row_idx, np.arange(W), indexing="ij")
    xs = (ii.ravel() + 0.5 - W / 2) / focal
    ys = (jj.ravel() + 0.5 - H / 2) / focal
    d = xs[:, None] * right + (-ys)[:, None] * upv + fwd
    d /= np.linalg.norm(d, axis=-1, keepdims=True)
    o = np.broadcast_to(eye, d.shape)
    return (torch.from_numpy(np.ascontiguousarray(o, np.float32)).to(device), torch.from_numpy(np.ascontiguousarray(d, np.float32)).to(device))


def occupancy_bitfield(device, H=128, bound=1.0):
    """Analytic occupancy of the scene at cell centres, dilated by one cell, packed in Morton order (cascade 1)."""
    r = torch.arange(H, dtype=torch.int32, device=device)
    xx, yy, zz = torch.meshgrid(r, r, r, indexing="ij")
    coords = torch.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], -1).contiguous()
    centres = (coords.float() + 0.5) / H * 2 - 1
    occ = (density(centres * bound) > 0).float().view(1, 1, H, H, H)
    occ = torch.nn.functional.max_pool3d(occ, 3, stride=1, padding=1).view(-1)
    idx = _raymarching.morton3D(coords).long()
    grid = torch.zeros(H ** 3, device=device)
    grid[idx] = occ
    return _raymarching.packbits(grid.view(1, -1), 0.5)


@torch.no_grad()
def render_ground_truth(rays_o, rays_d, bound=1.0, bg_color=1.0, H=128, max_steps=1024, min_near=0.2, chunk=1 << 16):
    """Pixels of the analytic scene along the given rays, rendered with the same quadrature as training."""
    dev = rays_o.device
    bits = occupancy_bitfield(dev, H, bound)
    aabb = torch.tensor([-bound, -bound, -bound, bound, bound, bound], dtype=torch.float32, device=dev)
    out = torch.empty(rays_o.shape[0], 3, device=dev)
    for a in range(0, rays_o.shape[0], chunk):
        o, d = rays_o[a:a + chunk].contiguous(), rays_d[a:a + chunk].contiguous()
        nears, fars = _raymarching.near_far_from_aabb(o, d, aabb, min_near)
        xyzs, dirs, deltas, rays = _raymarching.march_rays_train(o, d, bound, bits, 1, H, nears, fars, None, -1, False, 128, True, 0.0, max_steps)
        sig, rgb = density(xyzs), colour(xyzs).contiguous()
        ws, depth, image = _raymarching.composite_rays_train(sig, rgb, deltas, rays, 1e-4)
        out[a:a + chunk] = image + (1 - ws[:, None]) * bg_color
    return out


def make_pool(n_views, rays_per_view, device, seed=0, bound=1.0, bg_color=1.0):
    o, d = random_rays(n_views, rays_per_view, device, seed)
    gt = render_ground_truth(o, d, bound, bg_color)
    perm = torch.randperm(o.shape[0], generator=torch.Generator().manual_seed(seed + 1)).to(device)
    return o[perm].contiguous(), d[perm].contiguous(), gt[perm].contiguous()
