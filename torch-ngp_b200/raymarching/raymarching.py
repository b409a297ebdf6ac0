"""raymarching ops -- same public API as /root/reference/raymarching/raymarching.py
(near_far_from_aabb :19-49, sph_from_ray :52-80, morton3D :83-104, morton3D_invert :106-126,
packbits :129-155, march_rays_train :161-235, composite_rays_train :238-292, march_rays :298-349,
composite_rays :352-374), backed by libngp_b200.so.  Signatures, defaults, returned shapes and
the fp32 casting under autocast are the reference's.
"""
import torch
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

from . import backend as _backend


def _to_cuda(t):
    return t if t.is_cuda else t.cuda()


class _near_far_from_aabb(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, rays_o, rays_d, aabb, min_near=0.2):
        """rays_o/d [N,3], aabb [6] (xmin,ymin,zmin,xmax,ymax,zmax) -> nears [N], fars [N]."""
        rays_o = _to_cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _to_cuda(rays_d).contiguous().view(-1, 3)
        aabb = _to_cuda(aabb).contiguous()
        N = rays_o.shape[0]
        nears = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
        fars = torch.empty(N, dtype=rays_o.dtype, device=rays_o.device)
        _backend.near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars)
        return nears, fars


near_far_from_aabb = _near_far_from_aabb.apply


class _sph_from_ray(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, rays_o, rays_d, radius):
        """Background-sphere coordinates (theta, phi in [-1,1]) of each ray's far intersection."""
        rays_o = _to_cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _to_cuda(rays_d).contiguous().view(-1, 3)
        N = rays_o.shape[0]
        coords = torch.empty(N, 2, dtype=rays_o.dtype, device=rays_o.device)
        _backend.sph_from_ray(rays_o, rays_d, radius, N, coords)
        return coords


sph_from_ray = _sph_from_ray.apply


class _morton3D(Function):
    @staticmethod
    def forward(ctx, coords):
        """coords [N,3] int32 in [0,1024) -> Morton indices [N] int32."""
        coords = _to_cuda(coords)
        N = coords.shape[0]
        indices = torch.empty(N, dtype=torch.int32, device=coords.device)
        _backend.morton3D(coords.int().contiguous(), N, indices)
        return indices


morton3D = _morton3D.apply


class _morton3D_invert(Function):
    @staticmethod
    def forward(ctx, indices):
        """indices [N] int32 -> coords [N,3] int32."""
        indices = _to_cuda(indices)
        N = indices.shape[0]
        coords = torch.empty(N, 3, dtype=torch.int32, device=indices.device)
        _backend.morton3D_invert(indices.int().contiguous(), N, coords)
        return coords


morton3D_invert = _morton3D_invert.apply


class _packbits(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, grid, thresh, bitfield=None):
        """grid [C, H^3] float -> bitfield [C*H^3/8] uint8, bit i of byte n = (grid[8n+i] > thresh)."""
        grid = _to_cuda(grid).contiguous()
        C, H3 = grid.shape[0], grid.shape[1]
        N = C * H3 // 8
        if bitfield is None:
            bitfield = torch.empty(N, dtype=torch.uint8, device=grid.device)
        _backend.packbits(grid, N, thresh, bitfield)
        return bitfield


packbits = _packbits.apply


class _march_rays_train(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, rays_o, rays_d, bound, density_bitfield, C, H, nears, fars, step_counter=None, mean_count=-1,
                perturb=False, align=-1, force_all_rays=False, dt_gamma=0, max_steps=1024):
        """-> xyzs [M,3], dirs [M,3], deltas [M,2], rays [N,3] = (ray id, sample offset, sample count).

        Same sizing rules as the reference (:196-231): M = N*max_steps until a positive mean_count is
        supplied, then mean_count rounded up past the next multiple of `align`; when the full-size
        buffer is used the sample count is read back and the outputs are trimmed.
        Placement differs by design: rays[n] describes ray n and offsets follow ray order
        (the reference's order is decided by atomicAdd arrival, raymarching.cu:405-406).
        """
        rays_o = _to_cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _to_cuda(rays_d).contiguous().view(-1, 3)
        density_bitfield = _to_cuda(density_bitfield).contiguous()
        N = rays_o.shape[0]
        M = N * max_steps
        if not force_all_rays and mean_count > 0:
            if align > 0:
                mean_count += align - mean_count % align
            M = mean_count

        xyzs = torch.zeros(M, 3, dtype=rays_o.dtype, device=rays_o.device)
        dirs = torch.zeros(M, 3, dtype=rays_o.dtype, device=rays_o.device)
        deltas = torch.zeros(M, 2, dtype=rays_o.dtype, device=rays_o.device)
        rays = torch.empty(N, 3, dtype=torch.int32, device=rays_o.device)
        if step_counter is None:
            step_counter = torch.zeros(2, dtype=torch.int32, device=rays_o.device)
        if perturb:
            noises = torch.rand(N, dtype=rays_o.dtype, device=rays_o.device)
        else:
            noises = torch.zeros(N, dtype=rays_o.dtype, device=rays_o.device)

        _backend.march_rays_train(rays_o, rays_d, density_bitfield, bound, dt_gamma, max_steps, N, C, H, M, nears.contiguous(),
                                  fars.contiguous(), xyzs, dirs, deltas, rays, step_counter, noises)

        if force_all_rays or mean_count <= 0:
            m = step_counter[0].item()  # D2H, first steps only (as the reference)
            if align > 0:
                m += align - m % align
            xyzs, dirs, deltas = xyzs[:m], dirs[:m], deltas[:m]
        return xyzs, dirs, deltas, rays


march_rays_train = _march_rays_train.apply


class _composite_rays_train(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, sigmas, rgbs, deltas, rays, T_thresh=1e-4):
        """sigmas [M], rgbs [M,3], deltas [M,2], rays [N,3] -> weights_sum [N], depth [N], image [N,3]."""
        sigmas = sigmas.contiguous()
        rgbs = rgbs.contiguous()
        deltas = deltas.contiguous()
        M, N = sigmas.shape[0], rays.shape[0]
        weights_sum = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        depth = torch.empty(N, dtype=sigmas.dtype, device=sigmas.device)
        image = torch.empty(N, 3, dtype=sigmas.dtype, device=sigmas.device)
        _backend.composite_rays_train_forward(sigmas, rgbs, deltas, rays, M, N, T_thresh, weights_sum, depth, image)
        ctx.save_for_backward(sigmas, rgbs, deltas, rays, weights_sum, depth, image)
        ctx.dims = [M, N, T_thresh]
        return weights_sum, depth, image

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad_weights_sum, grad_depth, grad_image):
        grad_weights_sum = grad_weights_sum.contiguous()
        grad_depth = grad_depth.contiguous()
        grad_image = grad_image.contiguous()
        sigmas, rgbs, deltas, rays, weights_sum, depth, image = ctx.saved_tensors
        M, N, T_thresh = ctx.dims
        grad_sigmas = torch.zeros_like(sigmas)
        grad_rgbs = torch.zeros_like(rgbs)
        _backend.composite_rays_train_backward(grad_weights_sum, grad_depth, grad_image, sigmas, rgbs, deltas, rays, weights_sum,
                                               depth, image, M, N, T_thresh, grad_sigmas, grad_rgbs)
        return grad_sigmas, grad_rgbs, None, None, None


composite_rays_train = _composite_rays_train.apply


class _march_rays(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, density_bitfield, C, H, near, far, align=-1,
                perturb=False, dt_gamma=0, max_steps=1024):
        """Inference marcher: up to n_step samples for each of the first n_alive ids of rays_alive.
        -> xyzs [M',3], dirs [M',3], deltas [M',2], M' = n_alive*n_step rounded up past `align`."""
        rays_o = _to_cuda(rays_o).contiguous().view(-1, 3)
        rays_d = _to_cuda(rays_d).contiguous().view(-1, 3)
        M = n_alive * n_step
        if align > 0:
            M += align - (M % align)
        xyzs = torch.zeros(M, 3, dtype=rays_o.dtype, device=rays_o.device)
        dirs = torch.zeros(M, 3, dtype=rays_o.dtype, device=rays_o.device)
        deltas = torch.zeros(M, 2, dtype=rays_o.dtype, device=rays_o.device)
        if perturb:
            noises = torch.rand(n_alive, dtype=rays_o.dtype, device=rays_o.device)
        else:
            noises = torch.zeros(n_alive, dtype=rays_o.dtype, device=rays_o.device)
        _backend.march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, dt_gamma, max_steps, C, H,
                            density_bitfield, near, far, xyzs, dirs, deltas, noises)
        return xyzs, dirs, deltas


march_rays = _march_rays.apply


class _composite_rays(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)
    def forward(ctx, n_alive, n_step, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image, T_thresh=1e-2):
        """In place on weights_sum / depth / image / rays_t; finished rays get rays_alive[n] = -1."""
        _backend.composite_rays(n_alive, n_step, T_thresh, rays_alive, rays_t, sigmas.contiguous(), rgbs.contiguous(),
                                deltas.contiguous(), weights_sum, depth, image)
        return tuple()


composite_rays = _composite_rays.apply


def mark_untrained_grid(poses, intrinsic, density_grid, bound, return_count=False):
    """NeRFRenderer.mark_untrained_grid (nerf/renderer.py:433-496) on a density grid [cascade, H^3] (Morton order), in
    place: cells no training camera sees become -1.  poses [B,4,4] camera-to-world (tensor or ndarray), intrinsic
    (fx, fy, cx, cy).  One kernel instead of the reference's loops over 64^3 cell blocks x cascades x 64-camera batches."""
    import numpy as np
    if isinstance(poses, np.ndarray):
        poses = torch.from_numpy(poses)
    poses = poses.to(density_grid.device, torch.float32).contiguous()
    C, H3 = density_grid.shape
    H = round(H3 ** (1.0 / 3.0))
    assert H * H * H == H3, "density_grid must be [cascade, H^3]"
    fx, fy, cx, cy = [float(v) for v in intrinsic]
    count = torch.empty(C, H3, dtype=torch.int32, device=density_grid.device) if return_count else None
    _backend.mark_untrained_grid(poses, fx, fy, cx, cy, bound, C, H, density_grid, count)
    return count


def compact_alive(rays_alive, n_alive=None):
    """Device-side `rays_alive[rays_alive >= 0]` (nerf/renderer.py:416) without a boolean-mask sync.
    -> (compacted ids in a same-sized buffer, device int32[1] count)."""
    n = rays_alive.shape[0] if n_alive is None else n_alive
    out = torch.empty_like(rays_alive)
    n_out = torch.zeros(1, dtype=torch.int32, device=rays_alive.device)
    _backend.compact_alive(rays_alive, n, out, n_out)
    return out, n_out
