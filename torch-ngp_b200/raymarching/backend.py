"""`_backend` shim: the reference's pybind11 module `_raymarching` (raymarching/src/bindings.cpp:5-18),
same ten names and the same positional arguments, forwarding to the C ABI of libngp_b200.so.

Unlike the reference (whose host functions check nothing, raymarching.cu:148-156,482-490) every
tensor is validated: CUDA, contiguous, dtype, minimum size."""
import torch

import ngp_abi as abi

_f32, _i32, _u8 = torch.float32, torch.int32, torch.uint8
_ws_cache = {}


def _workspace(device, nbytes):
    """Small per-device scratch (block totals of the scans); grown on demand, never shrunk."""
    key = (device.type, device.index)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def near_far_from_aabb(rays_o, rays_d, aabb, N, min_near, nears, fars):
    abi.check_tensor(rays_o, "rays_o", _f32, "a float", 3 * N)
    abi.check_tensor(rays_d, "rays_d", _f32, "a float", 3 * N)
    abi.check_tensor(aabb, "aabb", _f32, "a float", 6)
    abi.check_tensor(nears, "nears", _f32, "a float", N)
    abi.check_tensor(fars, "fars", _f32, "a float", N)
    abi.check(abi.lib.ngp_near_far_from_aabb(rays_o.data_ptr(), rays_d.data_ptr(), aabb.data_ptr(), N, float(min_near),
                                             nears.data_ptr(), fars.data_ptr(), abi.stream()))


def sph_from_ray(rays_o, rays_d, radius, N, coords):
    abi.check_tensor(rays_o, "rays_o", _f32, "a float", 3 * N)
    abi.check_tensor(rays_d, "rays_d", _f32, "a float", 3 * N)
    abi.check_tensor(coords, "coords", _f32, "a float", 2 * N)
    abi.check(abi.lib.ngp_sph_from_ray(rays_o.data_ptr(), rays_d.data_ptr(), float(radius), N, coords.data_ptr(), abi.stream()))


def morton3D(coords, N, indices):
    abi.check_tensor(coords, "coords", _i32, "an int", 3 * N)
    abi.check_tensor(indices, "indices", _i32, "an int", N)
    abi.check(abi.lib.ngp_morton3D(coords.data_ptr(), N, indices.data_ptr(), abi.stream()))


def morton3D_invert(indices, N, coords):
    abi.check_tensor(indices, "indices", _i32, "an int", N)
    abi.check_tensor(coords, "coords", _i32, "an int", 3 * N)
    abi.check(abi.lib.ngp_morton3D_invert(indices.data_ptr(), N, coords.data_ptr(), abi.stream()))


def mark_untrained_grid(poses, fx, fy, cx, cy, bound, C, H, density_grid, count=None):
    B = poses.shape[0]
    abi.check_tensor(poses, "poses", _f32, "a float", 16 * B)
    abi.check_tensor(density_grid, "density_grid", _f32, "a float", C * H * H * H)
    if count is not None:
        abi.check_tensor(count, "count", torch.int32, "an int", C * H * H * H)
    abi.check(abi.lib.ngp_mark_untrained_grid(poses.data_ptr(), B, float(fx), float(fy), float(cx), float(cy), float(bound), C, H,
                                              density_grid.data_ptr(), abi.ptr(count), abi.stream()))


def packbits(grid, N, density_thresh, bitfield):
    abi.check_tensor(grid, "grid", _f32, "a float", 8 * N)
    abi.check_tensor(bitfield, "bitfield", _u8, "a uint8", N)
    abi.check(abi.lib.ngp_packbits(grid.data_ptr(), N, float(density_thresh), bitfield.data_ptr(), abi.stream()))


def march_rays_train(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears, fars, xyzs, dirs, deltas, rays,
                     counter, noises):
    abi.check_tensor(rays_o, "rays_o", _f32, "a float", 3 * N)
    abi.check_tensor(rays_d, "rays_d", _f32, "a float", 3 * N)
    abi.check_tensor(grid, "grid", _u8, "a uint8", C * H * H * H // 8)
    abi.check_tensor(nears, "nears", _f32, "a float", N)
    abi.check_tensor(fars, "fars", _f32, "a float", N)
    abi.check_tensor(xyzs, "xyzs", _f32, "a float", 3 * M)
    abi.check_tensor(dirs, "dirs", _f32, "a float", 3 * M)
    abi.check_tensor(deltas, "deltas", _f32, "a float", 2 * M)
    abi.check_tensor(rays, "rays", _i32, "an int", 3 * N)
    abi.check_tensor(counter, "counter", _i32, "an int", 2)
    abi.check_tensor(noises, "noises", _f32, "a float", N)
    # the larger workspace also carries the emit masks of pass 1 (pass 2 replays them); fall back to the scan-only one
    # when that would be unreasonably large (many rays x large bound)
    nbytes = abi.lib.ngp_march_rays_train_workspace_bytes_ex(N, float(bound), max_steps)
    if nbytes > (256 << 20):
        nbytes = abi.lib.ngp_march_rays_train_workspace_bytes(N)
    ws = _workspace(rays_o.device, nbytes)
    abi.check(abi.lib.ngp_march_rays_train(
        rays_o.data_ptr(), rays_d.data_ptr(), grid.data_ptr(), float(bound), float(dt_gamma), max_steps, N, C, H, M,
        nears.data_ptr(), fars.data_ptr(), xyzs.data_ptr(), dirs.data_ptr(), deltas.data_ptr(), rays.data_ptr(),
        counter.data_ptr(), noises.data_ptr(), ws.data_ptr(), ws.numel(), abi.stream()))


def composite_rays_train_forward(sigmas, rgbs, deltas, rays, M, N, T_thresh, weights_sum, depth, image):
    abi.check_tensor(sigmas, "sigmas", _f32, "a float", M)
    abi.check_tensor(rgbs, "rgbs", _f32, "a float", 3 * M)
    abi.check_tensor(deltas, "deltas", _f32, "a float", 2 * M)
    abi.check_tensor(rays, "rays", _i32, "an int", 3 * N)
    abi.check_tensor(weights_sum, "weights_sum", _f32, "a float", N)
    abi.check_tensor(depth, "depth", _f32, "a float", N)
    abi.check_tensor(image, "image", _f32, "a float", 3 * N)
    abi.check(abi.lib.ngp_composite_rays_train_forward(sigmas.data_ptr(), rgbs.data_ptr(), deltas.data_ptr(), rays.data_ptr(), M, N,
                                                       float(T_thresh), weights_sum.data_ptr(), depth.data_ptr(), image.data_ptr(),
                                                       abi.stream()))


def composite_rays_train_backward(grad_weights_sum, grad_depth, grad_image, sigmas, rgbs, deltas, rays, weights_sum, depth, image,
                                  M, N, T_thresh, grad_sigmas, grad_rgbs):
    for t, name, n in ((grad_weights_sum, "grad_weights_sum", N), (grad_depth, "grad_depth", N), (grad_image, "grad_image", 3 * N),
                       (sigmas, "sigmas", M), (rgbs, "rgbs", 3 * M), (deltas, "deltas", 2 * M), (weights_sum, "weights_sum", N),
                       (depth, "depth", N), (image, "image", 3 * N), (grad_sigmas, "grad_sigmas", M), (grad_rgbs, "grad_rgbs", 3 * M)):
        abi.check_tensor(t, name, _f32, "a float", n)
    abi.check_tensor(rays, "rays", _i32, "an int", 3 * N)
    abi.check(abi.lib.ngp_composite_rays_train_backward(
        grad_weights_sum.data_ptr(), grad_depth.data_ptr(), grad_image.data_ptr(), sigmas.data_ptr(), rgbs.data_ptr(),
        deltas.data_ptr(), rays.data_ptr(), weights_sum.data_ptr(), depth.data_ptr(), image.data_ptr(), M, N, float(T_thresh),
        grad_sigmas.data_ptr(), grad_rgbs.data_ptr(), abi.stream()))


def march_rays(n_alive, n_step, rays_alive, rays_t, rays_o, rays_d, bound, dt_gamma, max_steps, C, H, grid, nears, fars, xyzs,
               dirs, deltas, noises, n_alive_dev=None):
    M = n_alive * n_step
    abi.check_tensor(rays_alive, "rays_alive", _i32, "an int", n_alive)
    abi.check_tensor(rays_t, "rays_t", _f32, "a float")
    abi.check_tensor(rays_o, "rays_o", _f32, "a float")
    abi.check_tensor(rays_d, "rays_d", _f32, "a float")
    abi.check_tensor(grid, "grid", _u8, "a uint8", C * H * H * H // 8)
    abi.check_tensor(nears, "nears", _f32, "a float")
    abi.check_tensor(fars, "fars", _f32, "a float")
    abi.check_tensor(xyzs, "xyzs", _f32, "a float", 3 * M)
    abi.check_tensor(dirs, "dirs", _f32, "a float", 3 * M)
    abi.check_tensor(deltas, "deltas", _f32, "a float", 2 * M)
    abi.check_tensor(noises, "noises", _f32, "a float", n_alive)
    abi.check(abi.lib.ngp_march_rays(n_alive, abi.ptr(n_alive_dev), n_step, rays_alive.data_ptr(), rays_t.data_ptr(), rays_o.data_ptr(),
                                     rays_d.data_ptr(), float(bound), float(dt_gamma), max_steps, C, H, grid.data_ptr(),
                                     nears.data_ptr(), fars.data_ptr(), xyzs.data_ptr(), dirs.data_ptr(), deltas.data_ptr(),
                                     noises.data_ptr(), abi.stream()))


def composite_rays(n_alive, n_step, T_thresh, rays_alive, rays_t, sigmas, rgbs, deltas, weights_sum, depth, image, n_alive_dev=None):
    M = n_alive * n_step
    abi.check_tensor(rays_alive, "rays_alive", _i32, "an int", n_alive)
    abi.check_tensor(rays_t, "rays_t", _f32, "a float")
    abi.check_tensor(sigmas, "sigmas", _f32, "a float", M)
    abi.check_tensor(rgbs, "rgbs", _f32, "a float", 3 * M)
    abi.check_tensor(deltas, "deltas", _f32, "a float", 2 * M)
    abi.check_tensor(weights_sum, "weights_sum", _f32, "a float")
    abi.check_tensor(depth, "depth", _f32, "a float")
    abi.check_tensor(image, "image", _f32, "a float")
    abi.check(abi.lib.ngp_composite_rays(n_alive, abi.ptr(n_alive_dev), n_step, float(T_thresh), rays_alive.data_ptr(), rays_t.data_ptr(),
                                         sigmas.data_ptr(), rgbs.data_ptr(), deltas.data_ptr(), weights_sum.data_ptr(),
                                         depth.data_ptr(), image.data_ptr(), abi.stream()))


def compact_alive(rays_alive_in, n, rays_alive_out, n_out, n_dev=None):
    abi.check_tensor(rays_alive_in, "rays_alive_in", _i32, "an int", n)
    abi.check_tensor(rays_alive_out, "rays_alive_out", _i32, "an int", n)
    abi.check_tensor(n_out, "n_out", _i32, "an int", 1)
    nbytes = abi.lib.ngp_compact_alive_workspace_bytes(n)
    ws = _workspace(rays_alive_in.device, nbytes)
    abi.check(abi.lib.ngp_compact_alive(rays_alive_in.data_ptr(), n, abi.ptr(n_dev), rays_alive_out.data_ptr(), n_out.data_ptr(),
                                        ws.data_ptr(), ws.numel(), abi.stream()))
