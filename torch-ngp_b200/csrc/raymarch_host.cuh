// raymarch_host.cuh -- C ABI entry points of the ray-marching ops.  Included by raymarch.cu only.
#pragma once

namespace ngp {
namespace {
inline size_t scan_ws_bytes(uint32_t nblocks) { return sizeof(uint32_t) * (2 + (size_t)(nblocks ? nblocks : 1)); }
inline bool aligned(const void *p, size_t a) { return (reinterpret_cast<uintptr_t>(p) & (a - 1)) == 0; }
// Emit masks of the warp-per-ray marcher: a ray visits at most (far - near) / dt_min + 1 <= bound * max_steps + 1 points
// (aabb diagonal 2*sqrt(3)*bound, dt_min = 2*sqrt(3)/max_steps); + slack for rounding, + the leading count word.
inline uint32_t mask_words_for(float bound, uint32_t max_steps) {
    const double pts = (double)(bound < 1.f ? 1.f : bound) * (double)max_steps + 8.0;
    return (uint32_t)((pts + 31.0) / 32.0) + 2u;
}
inline size_t mask_ws_offset(uint32_t N) { return (scan_ws_bytes(div_round_up(N, 8u)) + 15u) & ~(size_t)15u; }
}  // namespace
}  // namespace ngp

extern "C" {

int ngp_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, uint32_t N, float min_near,
                           float *nears, float *fars, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(rays_o && rays_d && aabb && nears && fars, "near_far_from_aabb: null pointer argument");
    { NGP_PROF("k_near_far", (cudaStream_t)stream); k_near_far<<<div_round_up(N, 256u), 256, 0, (cudaStream_t)stream>>>(rays_o, rays_d, aabb, N, min_near, nears, fars); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_sph_from_ray(const float *rays_o, const float *rays_d, float radius, uint32_t N, float *coords, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(rays_o && rays_d && coords, "sph_from_ray: null pointer argument");
    { NGP_PROF("k_sph_from_ray", (cudaStream_t)stream); k_sph_from_ray<<<div_round_up(N, 256u), 256, 0, (cudaStream_t)stream>>>(rays_o, rays_d, radius, N, coords); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_morton3D(const int *coords, uint32_t N, int *indices, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(coords && indices, "morton3D: null pointer argument");
    { NGP_PROF("k_morton3D", (cudaStream_t)stream); k_morton3D<<<div_round_up(N, 256u), 256, 0, (cudaStream_t)stream>>>(coords, N, indices); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_morton3D_invert(const int *indices, uint32_t N, int *coords, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(indices && coords, "morton3D_invert: null pointer argument");
    { NGP_PROF("k_morton3D_invert", (cudaStream_t)stream); k_morton3D_invert<<<div_round_up(N, 256u), 256, 0, (cudaStream_t)stream>>>(indices, N, coords); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_mark_untrained_grid(const float *poses, uint32_t B, double fx, double fy, double cx, double cy, float bound, uint32_t C, uint32_t H,
                            float *density_grid, int *count_out, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(density_grid && (B == 0 || poses), "mark_untrained_grid: null pointer argument");
    NGP_CHECK_ARG(C >= 1 && H >= 2 && H <= 1024 && fx != 0.0 && fy != 0.0, "mark_untrained_grid: bad cascade / grid size / focal length");
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t H3 = H * H * H;
    { NGP_PROF("k_mark_untrained", st); k_mark_untrained<<<dim3(div_round_up(H3, 256u), C, 1), 256, 0, st>>>(poses, B, (float)(cx / fx), (float)(cy / fy), bound, C, H, density_grid, count_out); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_packbits(const float *grid, uint32_t N, float density_thresh, uint8_t *bitfield, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(grid && bitfield, "packbits: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t done = 0;
    if (aligned(grid, 16) && aligned(bitfield, 4) && N >= 4) {
        const uint32_t N4 = N / 4;
        { NGP_PROF("k_packbits32", st); k_packbits32<<<div_round_up(N4, 256u), 256, 0, st>>>(reinterpret_cast<const float4 *>(grid), N4, density_thresh,
                                                             reinterpret_cast<uint32_t *>(bitfield)); }
        NGP_CHECK_LAUNCH();
        done = N4 * 4;
    }
    if (done < N) {
        { NGP_PROF("k_packbits_tail", st); k_packbits_tail<<<div_round_up(N - done, 256u), 256, 0, st>>>(grid, done, N, density_thresh, bitfield); }
        NGP_CHECK_LAUNCH();
    }
    return NGP_OK;
}

size_t ngp_march_rays_train_workspace_bytes(uint32_t N) {
    return ngp::scan_ws_bytes(ngp::div_round_up(N, (uint32_t)ngp::kWarpRays));  // the warp-per-ray variant has the most blocks
}

size_t ngp_march_rays_train_workspace_bytes_ex(uint32_t N, float bound, uint32_t max_steps) {
    return ngp::mask_ws_offset(N) + (size_t)N * ngp::mask_words_for(bound, max_steps) * sizeof(uint32_t);
}

int ngp_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                         uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float *nears,
                         const float *fars, float *xyzs, float *dirs, float *deltas, int *rays, int *counter,
                         const float *noises, void *workspace, size_t workspace_bytes, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(rays_o && rays_d && grid && nears && fars && xyzs && dirs && deltas && rays && counter && noises,
                  "march_rays_train: null pointer argument");
    NGP_CHECK_ARG(C >= 1 && C <= 16 && H >= 1 && H <= 1024, "march_rays_train: C=%u / H=%u out of range", C, H);
    NGP_CHECK_ARG(aligned(deltas, 8), "march_rays_train: deltas must be 8-byte aligned");
    const bool warp = get_option("march_variant", -1) != 0;  // -1 auto / 1: warp per ray; 0: thread per ray
    const uint32_t nb = div_round_up(N, warp ? (uint32_t)kWarpRays : (uint32_t)kMarchBlock);
    NGP_CHECK_ARG(workspace && workspace_bytes >= scan_ws_bytes(nb), "march_rays_train: workspace too small (%zu < %zu bytes)",
                  workspace_bytes, scan_ws_bytes(nb));
    cudaStream_t st = (cudaStream_t)stream;
    march_ws_t *ws = reinterpret_cast<march_ws_t *>(workspace);
    NGP_CHECK_CUDA(cudaMemsetAsync(ws, 0, 2 * sizeof(uint32_t), st));
    if (warp) {
        // a workspace sized by ngp_march_rays_train_workspace_bytes_ex also carries the per-ray emit masks: pass 2 then
        // replays them instead of repeating the occupancy tests
        const uint32_t mw = mask_words_for(bound, max_steps);
        uint32_t *masks = nullptr;
        if (workspace_bytes >= mask_ws_offset(N) + (size_t)N * mw * sizeof(uint32_t) && get_option("march_masks", 1))
            masks = reinterpret_cast<uint32_t *>(reinterpret_cast<uint8_t *>(workspace) + mask_ws_offset(N));
        { NGP_PROF("k_march_train_count_warp", st); k_march_train_count_warp<<<nb, kWarpRays * 32, 0, st>>>(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, nears, fars,
                                                                  noises, rays, counter, ws, masks, mw); }
        NGP_CHECK_LAUNCH();
        { NGP_PROF("k_march_train_write_warp", st); k_march_train_write_warp<<<nb, kWarpRays * 32, 0, st>>>(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears,
                                                                  fars, noises, xyzs, dirs, deltas, rays, ws, masks, mw); }
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    { NGP_PROF("k_march_train_count", st); k_march_train_count<<<nb, kMarchBlock, 0, st>>>(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, nears, fars, noises,
                                                   rays, counter, ws); }
    NGP_CHECK_LAUNCH();
    { NGP_PROF("k_march_train_write", st); k_march_train_write<<<nb, kMarchBlock, 0, st>>>(rays_o, rays_d, grid, bound, dt_gamma, max_steps, N, C, H, M, nears, fars, noises,
                                                   xyzs, dirs, deltas, rays, ws); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas, const int *rays,
                                     uint32_t M, uint32_t N, float T_thresh, float *weights_sum, float *depth,
                                     float *image, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(rays && weights_sum && depth && image, "composite_rays_train_forward: null pointer argument");
    NGP_CHECK_ARG(M == 0 || (sigmas && rgbs && deltas), "composite_rays_train_forward: null sample buffers with M=%u", M);
    if (get_option("composite_variant", -1) != 0) {
        { NGP_PROF("k_composite_train_fwd_warp", (cudaStream_t)stream); k_composite_train_fwd_warp<<<div_round_up(N, (uint32_t)kWarpRays), kWarpRays * 32, 0, (cudaStream_t)stream>>>(
            sigmas, rgbs, deltas, rays, M, N, T_thresh, weights_sum, depth, image); }
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    { NGP_PROF("k_composite_train_fwd", (cudaStream_t)stream); k_composite_train_fwd<<<div_round_up(N, 128u), 128, 0, (cudaStream_t)stream>>>(sigmas, rgbs, deltas, rays, M, N, T_thresh,
                                                                                 weights_sum, depth, image); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_depth, const float *grad_image,
                                      const float *sigmas, const float *rgbs, const float *deltas, const int *rays,
                                      const float *weights_sum, const float *depth, const float *image, uint32_t M,
                                      uint32_t N, float T_thresh, float *grad_sigmas, float *grad_rgbs, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(grad_weights_sum && grad_depth && grad_image && rays && weights_sum && depth && image,
                  "composite_rays_train_backward: null pointer argument");
    NGP_CHECK_ARG(M == 0 || (sigmas && rgbs && deltas && grad_sigmas && grad_rgbs),
                  "composite_rays_train_backward: null sample buffers with M=%u", M);
    if (get_option("composite_variant", -1) != 0) {
        { NGP_PROF("k_composite_train_bwd_warp", (cudaStream_t)stream); k_composite_train_bwd_warp<<<div_round_up(N, (uint32_t)kWarpRays), kWarpRays * 32, 0, (cudaStream_t)stream>>>(
            grad_weights_sum, grad_depth, grad_image, sigmas, rgbs, deltas, rays, weights_sum, depth, image, M, N, T_thresh, grad_sigmas, grad_rgbs); }
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    { NGP_PROF("k_composite_train_bwd", (cudaStream_t)stream); k_composite_train_bwd<<<div_round_up(N, 128u), 128, 0, (cudaStream_t)stream>>>(
        grad_weights_sum, grad_depth, grad_image, sigmas, rgbs, deltas, rays, weights_sum, depth, image, M, N, T_thresh,
        grad_sigmas, grad_rgbs); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_march_rays(uint32_t n_alive, const int *n_alive_dev, uint32_t n_step, const int *rays_alive, const float *rays_t,
                   const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                   uint32_t H, const uint8_t *grid, const float *nears, const float *fars, float *xyzs, float *dirs,
                   float *deltas, const float *noises, void *stream) {
    using namespace ngp;
    (void)nears;  // the reference reads it but never uses it (:746)
    NGP_CHECK_ARG(C >= 1 && C <= 16 && H >= 1 && H <= 1024, "march_rays: C=%u / H=%u out of range", C, H);
    if (n_alive == 0) return NGP_OK;
    NGP_CHECK_ARG(rays_alive && rays_t && rays_o && rays_d && grid && fars && xyzs && dirs && deltas && noises,
                  "march_rays: null pointer argument");
    NGP_CHECK_ARG(aligned(deltas, 8), "march_rays: deltas must be 8-byte aligned");
    // one warp per ray pays off once the launch no longer fills the GPU with threads (option = largest n_alive it is used for; 0 = never)
    if (n_alive <= (uint32_t)get_option("march_rays_warp", 32768)) {
        NGP_PROF("k_march_rays_warp", (cudaStream_t)stream);
        k_march_rays_warp<<<div_round_up(n_alive, (uint32_t)kWarpRays), kWarpRays * 32, 0, (cudaStream_t)stream>>>(
            n_alive, n_alive_dev, n_step, rays_alive, rays_t, rays_o, rays_d, bound, dt_gamma, max_steps, C, H, grid, fars, xyzs, dirs, deltas, noises);
    } else {
        NGP_PROF("k_march_rays", (cudaStream_t)stream);
        k_march_rays<<<div_round_up(n_alive, 128u), 128, 0, (cudaStream_t)stream>>>(n_alive, n_alive_dev, n_step, rays_alive, rays_t, rays_o,
                                                                                    rays_d, bound, dt_gamma, max_steps, C, H, grid, fars,
                                                                                    xyzs, dirs, deltas, noises);
    }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_composite_rays(uint32_t n_alive, const int *n_alive_dev, uint32_t n_step, float T_thresh, int *rays_alive,
                       float *rays_t, const float *sigmas, const float *rgbs, const float *deltas, float *weights_sum,
                       float *depth, float *image, void *stream) {
    using namespace ngp;
    if (n_alive == 0) return NGP_OK;
    NGP_CHECK_ARG(rays_alive && rays_t && sigmas && rgbs && deltas && weights_sum && depth && image,
                  "composite_rays: null pointer argument");
    NGP_CHECK_ARG(aligned(deltas, 8), "composite_rays: deltas must be 8-byte aligned");
    { NGP_PROF("k_composite_rays", (cudaStream_t)stream); k_composite_rays<<<div_round_up(n_alive, 128u), 128, 0, (cudaStream_t)stream>>>(n_alive, n_alive_dev, n_step, T_thresh, rays_alive,
                                                                                   rays_t, sigmas, rgbs, deltas, weights_sum, depth,
                                                                                   image); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

size_t ngp_compact_alive_workspace_bytes(uint32_t n) {
    return ngp::scan_ws_bytes(ngp::div_round_up(n, (uint32_t)(ngp::kMarchBlock * ngp::kCompactItems)));
}

int ngp_compact_alive(const int *in, uint32_t n, const int *n_dev, int *out, int *n_out, void *workspace,
                      size_t workspace_bytes, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(in && out && n_out && in != out, "compact_alive: null or aliasing in/out");
    const uint32_t nb = div_round_up(n ? n : 1u, (uint32_t)(kMarchBlock * kCompactItems));
    NGP_CHECK_ARG(workspace && workspace_bytes >= scan_ws_bytes(nb), "compact_alive: workspace too small");
    cudaStream_t st = (cudaStream_t)stream;
    march_ws_t *ws = reinterpret_cast<march_ws_t *>(workspace);
    NGP_CHECK_CUDA(cudaMemsetAsync(ws, 0, 2 * sizeof(uint32_t), st));
    { NGP_PROF("k_compact_count", st); k_compact_count<<<nb, kMarchBlock, 0, st>>>(in, n, n_dev, n_out, ws); }
    NGP_CHECK_LAUNCH();
    { NGP_PROF("k_compact_write", st); k_compact_write<<<nb, kMarchBlock, 0, st>>>(in, n, n_dev, out, ws); }
    NGP_CHECK_LAUNCH();
    { NGP_PROF("k_compact_publish", st); k_compact_publish<<<1, 1, 0, st>>>(ws, n_out); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

}  // extern "C"
