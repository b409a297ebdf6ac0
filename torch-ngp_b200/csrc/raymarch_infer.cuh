// raymarch_infer.cuh -- inference marcher / compositor (resumable, fixed slab per alive ray) and
// the device-side alive-ray compaction.  Included by raymarch.cu only.
#pragma once

namespace ngp {
namespace {

// reference :709-814.  n_alive_dev (optional) clamps n_alive from device memory so that a
// sync-free render loop can launch with an upper bound.
__global__ void __launch_bounds__(128) k_march_rays(uint32_t n_alive, const int *__restrict__ n_alive_dev, uint32_t n_step,
                                                    const int *__restrict__ rays_alive, const float *__restrict__ rays_t,
                                                    const float *__restrict__ rays_o, const float *__restrict__ rays_d, float bound,
                                                    float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H,
                                                    const uint8_t *__restrict__ grid, const float *__restrict__ fars,
                                                    float *__restrict__ xyzs, float *__restrict__ dirs, float *__restrict__ deltas,
                                                    const float *__restrict__ noises) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n_alive_dev) n_alive = min(n_alive, (uint32_t)max(*n_alive_dev, 0));
    if (n >= n_alive) return;
    const int index = rays_alive[n];
    const march_consts_t c = make_consts(bound, dt_gamma, max_steps, C, H, grid);
    const ray_t r = load_ray(rays_o, rays_d, (uint32_t)index);
    float *px = xyzs + 3 * (size_t)n * n_step, *pd = dirs + 3 * (size_t)n * n_step, *pl = deltas + 2 * (size_t)n * n_step;
    const float far = fars[index];
    float t = rays_t[index];
    t = __fmaf_rn(clampf(__fmul_rn(t, c.dt_gamma), c.dt_min, c.dt_max), noises[n], t);  // :755
    float last_t = t, x, y, z, dt;
    uint32_t step = 0;
    while (t < far && step < n_step) {
        if (probe(r, c, t, x, y, z, dt)) {
            px[0] = x; px[1] = y; px[2] = z;
            pd[0] = r.dx; pd[1] = r.dy; pd[2] = r.dz;
            t = __fadd_rn(t, dt);
            *reinterpret_cast<float2 *>(pl) = make_float2(dt, __fsub_rn(t, last_t));
            last_t = t;
            px += 3; pd += 3; pl += 2;
            step++;
        }
    }
}

// The same op with one WARP per alive ray (march_ray_warp of the training marcher: 32 consecutive t values tested in
// parallel, skip rule replayed on ballots -- samples bit-identical to the loop above).  A thread of the kernel above walks
// its ray alone, one dependent bitfield load per visited voxel; that is the whole cost of the first loop iteration of a
// frame (every ray crosses empty space to its first hit) and of the late iterations (few rays, up to 64 samples each).
__global__ void __launch_bounds__(kWarpRays * 32) k_march_rays_warp(uint32_t n_alive, const int *__restrict__ n_alive_dev, uint32_t n_step,
                                                                     const int *__restrict__ rays_alive, const float *__restrict__ rays_t,
                                                                     const float *__restrict__ rays_o, const float *__restrict__ rays_d, float bound,
                                                                     float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H,
                                                                     const uint8_t *__restrict__ grid, const float *__restrict__ fars,
                                                                     float *__restrict__ xyzs, float *__restrict__ dirs, float *__restrict__ deltas,
                                                                     const float *__restrict__ noises) {
    const uint32_t n = blockIdx.x * kWarpRays + (threadIdx.x >> 5);
    if (n_alive_dev) n_alive = min(n_alive, (uint32_t)max(*n_alive_dev, 0));
    if (n >= n_alive) return;  // whole warps leave together
    const int index = rays_alive[n];
    const march_consts_t c = make_consts(bound, dt_gamma, max_steps, C, H, grid);
    const ray_t r = load_ray(rays_o, rays_d, (uint32_t)index);
    float t = rays_t[index];
    t = __fmaf_rn(clampf(__fmul_rn(t, c.dt_gamma), c.dt_min, c.dt_max), noises[n], t);  // :755
    march_ray_warp<true>(r, c, t, fars[index], n_step, xyzs + 3 * (size_t)n * n_step, dirs + 3 * (size_t)n * n_step, deltas + 2 * (size_t)n * n_step);
}

// reference :827-914
__global__ void __launch_bounds__(128) k_composite_rays(uint32_t n_alive, const int *__restrict__ n_alive_dev, uint32_t n_step,
                                                        float T_thresh, int *__restrict__ rays_alive, float *__restrict__ rays_t,
                                                        const float *__restrict__ sigmas, const float *__restrict__ rgbs,
                                                        const float *__restrict__ deltas, float *__restrict__ weights_sum,
                                                        float *__restrict__ depth, float *__restrict__ image) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n_alive_dev) n_alive = min(n_alive, (uint32_t)max(*n_alive_dev, 0));
    if (n >= n_alive) return;
    const int index = rays_alive[n];
    const float *ps = sigmas + (size_t)n * n_step, *pc = rgbs + 3 * (size_t)n * n_step, *pl = deltas + 2 * (size_t)n * n_step;
    float t = rays_t[index], weight_sum = weights_sum[index], d = depth[index];
    float r = image[3 * (size_t)index], g = image[3 * (size_t)index + 1], b = image[3 * (size_t)index + 2];
    uint32_t step = 0;
    while (step < n_step) {
        const float2 dl = *reinterpret_cast<const float2 *>(pl + 2 * (size_t)step);
        if (dl.x == 0) break;  // the marcher left this slot untouched: ray finished
        const float alpha = 1.0f - __expf(-ps[step] * dl.x);
        const float T = 1 - weight_sum;  // :877
        const float weight = alpha * T;
        weight_sum += weight;
        t += dl.y;
        d += weight * t;
        r += weight * pc[3 * (size_t)step]; g += weight * pc[3 * (size_t)step + 1]; b += weight * pc[3 * (size_t)step + 2];
        if (T < T_thresh) break;
        step++;
    }
    if (step < n_step) rays_alive[n] = -1;
    else rays_t[index] = t;
    weights_sum[index] = weight_sum;
    depth[index] = d;
    image[3 * (size_t)index] = r; image[3 * (size_t)index + 1] = g; image[3 * (size_t)index + 2] = b;
}

// ------------------------------------------------------------------------------------------
// stable compaction of the non-negative ids (replaces `rays_alive[rays_alive >= 0]`,
// nerf/renderer.py:416, which costs a host sync per loop iteration in the reference)
// ------------------------------------------------------------------------------------------
constexpr int kCompactItems = 8;

__global__ void __launch_bounds__(kMarchBlock) k_compact_count(const int *__restrict__ in, uint32_t n, const int *__restrict__ n_dev,
                                                               int *__restrict__ n_out, march_ws_t *__restrict__ ws) {
    __shared__ uint32_t s_warp[kMarchBlock / 32];
    __shared__ uint32_t s_flag;
    if (n_dev) n = min(n, (uint32_t)max(*n_dev, 0));
    const uint32_t base = (blockIdx.x * kMarchBlock + threadIdx.x) * kCompactItems;
    uint32_t cnt = 0;
#pragma unroll
    for (int i = 0; i < kCompactItems; ++i)
        if (base + i < n && in[base + i] >= 0) cnt++;
    uint32_t total;
    block_excl_scan(cnt, s_warp, total);
    if (threadIdx.x == 0) {
        ws->block_sums[blockIdx.x] = total;
        __threadfence();
        s_flag = (atomicAdd(&ws->done, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_flag) return;
    __threadfence();
    const uint32_t nb = gridDim.x;
    const uint32_t chunk = div_round_up(nb, (uint32_t)kMarchBlock);
    const uint32_t lo = min(threadIdx.x * chunk, nb), hi = min(lo + chunk, nb);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += __ldcg(ws->block_sums + i);
    uint32_t grand;
    uint32_t run = block_excl_scan(sum, s_warp, grand);
    for (uint32_t i = lo; i < hi; ++i) {
        const uint32_t v = __ldcg(ws->block_sums + i);
        ws->block_sums[i] = run;
        run += v;
    }
    if (threadIdx.x == 0) {
        ws->pad = grand;  // k_compact_write publishes it to n_out (n_out may alias n_dev)
        ws->done = 0;
    }
    (void)n_out;
}

__global__ void __launch_bounds__(kMarchBlock) k_compact_write(const int *__restrict__ in, uint32_t n, const int *__restrict__ n_dev,
                                                               int *__restrict__ out, const march_ws_t *__restrict__ ws) {
    __shared__ uint32_t s_warp[kMarchBlock / 32];
    if (n_dev) n = min(n, (uint32_t)max(*n_dev, 0));
    const uint32_t base = (blockIdx.x * kMarchBlock + threadIdx.x) * kCompactItems;
    int v[kCompactItems];
    uint32_t cnt = 0;
#pragma unroll
    for (int i = 0; i < kCompactItems; ++i) {
        v[i] = (base + i < n) ? in[base + i] : -1;
        cnt += v[i] >= 0;
    }
    uint32_t total;
    uint32_t pos = block_excl_scan(cnt, s_warp, total) + ws->block_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < kCompactItems; ++i)
        if (v[i] >= 0) out[pos++] = v[i];
}

__global__ void k_compact_publish(const march_ws_t *__restrict__ ws, int *__restrict__ n_out) { *n_out = (int)ws->pad; }

}  // namespace
}  // namespace ngp
