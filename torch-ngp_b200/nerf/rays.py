"""get_rays of the reference (nerf/utils.py:52-136) with the ray construction on the device in one kernel.

Index selection keeps the reference's semantics and RNG calls (torch.randint / torch.multinomial on the device: uniform,
patch-based, or error-map weighted); what follows -- meshgrid, gather, normalise, rotate, expand: eight torch kernels and
a [B, H*W] pixel grid per call in the reference -- is ngp_get_rays."""
import torch

import ngp_abi as abi


@torch.no_grad()
def get_rays(poses, intrinsics, H, W, N=-1, error_map=None, patch_size=1):
    """poses [B,4,4] cam2world, intrinsics (fx, fy, cx, cy) -> dict(rays_o [B,N,3], rays_d [B,N,3], inds [B,N] if N > 0,
    inds_coarse with an error map) exactly as the reference returns them."""
    device = poses.device
    if not poses.is_cuda:
        raise RuntimeError("get_rays needs CUDA tensors (there is no CPU path)")
    poses = poses.to(torch.float32).contiguous()
    B = poses.shape[0]
    fx, fy, cx, cy = [float(v) for v in intrinsics]
    results = {}
    inds = None
    if N > 0:
        N = min(N, H * W)
        if patch_size > 1:
            num_patch = N // (patch_size ** 2)
            inds_x = torch.randint(0, H - patch_size, size=[num_patch], device=device)
            inds_y = torch.randint(0, W - patch_size, size=[num_patch], device=device)
            inds = torch.stack([inds_x, inds_y], dim=-1)
            pi, pj = torch.meshgrid(torch.arange(patch_size, device=device), torch.arange(patch_size, device=device), indexing="ij")
            offsets = torch.stack([pi.reshape(-1), pj.reshape(-1)], dim=-1)
            inds = (inds.unsqueeze(1) + offsets.unsqueeze(0)).view(-1, 2)
            inds = inds[:, 0] * W + inds[:, 1]
            N = inds.shape[0]
            inds = inds.expand([B, N])
        elif error_map is None:
            inds = torch.randint(0, H * W, size=[N], device=device).expand([B, N])  # may duplicate
        else:
            inds_coarse = torch.multinomial(error_map.to(device), N, replacement=False)  # [B, N] in [0, 128*128)
            inds_x, inds_y = inds_coarse // 128, inds_coarse % 128
            sx, sy = H / 128, W / 128
            inds_x = (inds_x * sx + torch.rand(B, N, device=device) * sx).long().clamp(max=H - 1)
            inds_y = (inds_y * sy + torch.rand(B, N, device=device) * sy).long().clamp(max=W - 1)
            inds = inds_x * W + inds_y
            results["inds_coarse"] = inds_coarse
        results["inds"] = inds
        inds32 = inds.to(torch.int32).contiguous()  # materialises the expand: [B, N]
        n, ptr, stride = N, inds32.data_ptr(), N
    else:
        n, ptr, stride = H * W, None, 0
    rays_o = torch.empty(B, n, 3, device=device)
    rays_d = torch.empty(B, n, 3, device=device)
    abi.check(abi.lib.ngp_get_rays(poses.data_ptr(), B, fx, fy, cx, cy, W, ptr, stride, n, rays_o.data_ptr(), rays_d.data_ptr(), abi.stream()))
    results["rays_o"], results["rays_d"] = rays_o, rays_d
    return results
