// raymarch_train.cuh -- training marcher (count + scan, then write) and training compositor.
// Included by raymarch.cu only.
#pragma once

namespace ngp {
namespace {

// workspace: [0] done-block counter, [1] pad, [2 ...] one uint32 per pass-1 block
struct march_ws_t {
    uint32_t done;
    uint32_t pad;
    uint32_t block_sums[1];
};

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) >= o) v += u;
    }
    return v;
}

// exclusive scan over the block (blockDim.x == kMarchBlock); returns this thread's exclusive prefix
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t *warp_tot /*[kMarchBlock/32]*/, uint32_t &total) {
    const uint32_t incl = warp_incl_scan(v);
    const uint32_t lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 31) warp_tot[wid] = incl;
    __syncthreads();
    uint32_t base = 0;
    total = 0;
#pragma unroll
    for (int w = 0; w < kMarchBlock / 32; ++w) {
        const uint32_t tw = warp_tot[w];
        if ((uint32_t)w < wid) base += tw;
        total += tw;
    }
    __syncthreads();
    return base + incl - v;
}

__device__ __forceinline__ float start_t(float near, const march_consts_t &c, float noise) {
    // t0 = near + clamp(near * dt_gamma, dt_min, dt_max) * noise      (reference :348-351, fused)
    return __fmaf_rn(clampf(__fmul_rn(near, c.dt_gamma), c.dt_min, c.dt_max), noise, near);
}

// pass 1: count occupied steps per ray, block-scan, publish the block total; the last block to
// arrive turns the block totals into exclusive prefixes (plus counter[0]) and bumps the counters.
__global__ void __launch_bounds__(kMarchBlock) k_march_train_count(
    const float *__restrict__ rays_o, const float *__restrict__ rays_d, const uint8_t *__restrict__ grid, float bound,
    float dt_gamma, uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, const float *__restrict__ nears,
    const float *__restrict__ fars, const float *__restrict__ noises, int *__restrict__ rays, int *__restrict__ counter,
    march_ws_t *__restrict__ ws) {
    __shared__ uint32_t s_warp[kMarchBlock / 32];
    __shared__ uint32_t s_flag;
    const uint32_t n = blockIdx.x * kMarchBlock + threadIdx.x;
    uint32_t num_steps = 0;
    if (n < N) {
        const march_consts_t c = make_consts(bound, dt_gamma, max_steps, C, H, grid);
        const ray_t r = load_ray(rays_o, rays_d, n);
        const float far = fars[n];
        float t = start_t(nears[n], c, noises[n]), x, y, z, dt;
        while (t < far && num_steps < max_steps) {
            if (probe(r, c, t, x, y, z, dt)) {
                num_steps++;
                t = __fadd_rn(t, dt);
            }
        }
    }
    uint32_t total;
    const uint32_t local = block_excl_scan(num_steps, s_warp, total);
    if (n < N) {
        rays[3 * (size_t)n] = (int)n;
        rays[3 * (size_t)n + 1] = (int)local;  // block-local for now; pass 2 adds the block prefix
        rays[3 * (size_t)n + 2] = (int)num_steps;
    }
    if (threadIdx.x == 0) {
        ws->block_sums[blockIdx.x] = total;
        __threadfence();
        s_flag = (atomicAdd(&ws->done, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_flag) return;

    // last block: exclusive scan of gridDim.x block totals, kMarchBlock threads, contiguous chunks
    __threadfence();
    const uint32_t nb = gridDim.x;
    const uint32_t chunk = div_round_up(nb, (uint32_t)kMarchBlock);
    const uint32_t lo = min(threadIdx.x * chunk, nb), hi = min(lo + chunk, nb);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += __ldcg(ws->block_sums + i);
    uint32_t grand;
    const uint32_t base = (uint32_t)counter[0];  // read by every thread BEFORE thread 0 bumps it below
    uint32_t run = block_excl_scan(sum, s_warp, grand) + base;
    __syncthreads();  // as in the warp variant (raymarch_warp.cuh): no thread may still be reading counter[0] when it is updated
    for (uint32_t i = lo; i < hi; ++i) {
        const uint32_t v = __ldcg(ws->block_sums + i);
        ws->block_sums[i] = run;
        run += v;
    }
    if (threadIdx.x == 0) {
        counter[0] += (int)grand;  // reference: atomicAdd(counter, num_steps) per ray (:405)
        counter[1] += (int)N;      // reference: atomicAdd(counter + 1, 1) per ray (:406)
        ws->done = 0;
    }
}

// pass 2: re-march and write the samples (reference :404-479)
__global__ void __launch_bounds__(kMarchBlock) k_march_train_write(
    const float *__restrict__ rays_o, const float *__restrict__ rays_d, const uint8_t *__restrict__ grid, float bound,
    float dt_gamma, uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float *__restrict__ nears,
    const float *__restrict__ fars, const float *__restrict__ noises, float *__restrict__ xyzs, float *__restrict__ dirs,
    float *__restrict__ deltas, int *__restrict__ rays, const march_ws_t *__restrict__ ws) {
    const uint32_t n = blockIdx.x * kMarchBlock + threadIdx.x;
    if (n >= N) return;
    const uint32_t num_steps = (uint32_t)rays[3 * (size_t)n + 2];
    const uint32_t point_index = (uint32_t)rays[3 * (size_t)n + 1] + ws->block_sums[blockIdx.x];
    rays[3 * (size_t)n + 1] = (int)point_index;
    if (num_steps == 0) return;
    if (point_index + num_steps > M) return;  // dropped ray, samples stay as the caller initialised them (:416)

    const march_consts_t c = make_consts(bound, dt_gamma, max_steps, C, H, grid);
    const ray_t r = load_ray(rays_o, rays_d, n);
    const float far = fars[n];
    float t = start_t(nears[n], c, noises[n]), x, y, z, dt;
    float last_t = t;
    float *px = xyzs + 3 * (size_t)point_index, *pd = dirs + 3 * (size_t)point_index, *pl = deltas + 2 * (size_t)point_index;
    uint32_t step = 0;
    while (t < far && step < num_steps) {
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

// front-to-back compositing over each ray's segment (reference :500-577)
__global__ void __launch_bounds__(128) k_composite_train_fwd(const float *__restrict__ sigmas, const float *__restrict__ rgbs,
                                                             const float *__restrict__ deltas, const int *__restrict__ rays,
                                                             uint32_t M, uint32_t N, float T_thresh, float *__restrict__ weights_sum,
                                                             float *__restrict__ depth, float *__restrict__ image) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    const uint32_t index = (uint32_t)rays[3 * (size_t)n], offset = (uint32_t)rays[3 * (size_t)n + 1], num_steps = (uint32_t)rays[3 * (size_t)n + 2];
    float T = 1.0f, r = 0, g = 0, b = 0, ws = 0, t = 0, d = 0;
    if (!(num_steps == 0 || offset + num_steps > M)) {
        const float *ps = sigmas + offset, *pc = rgbs + 3 * (size_t)offset, *pl = deltas + 2 * (size_t)offset;
        for (uint32_t step = 0; step < num_steps; ++step) {
            const float2 dl = *reinterpret_cast<const float2 *>(pl + 2 * (size_t)step);
            const float alpha = 1.0f - __expf(-ps[step] * dl.x);
            const float weight = alpha * T;
            r += weight * pc[3 * (size_t)step]; g += weight * pc[3 * (size_t)step + 1]; b += weight * pc[3 * (size_t)step + 2];
            t += dl.y;
            d += weight * t;
            ws += weight;
            T *= 1.0f - alpha;
            if (T < T_thresh) break;
        }
    }
    weights_sum[index] = ws;
    depth[index] = d;
    image[3 * (size_t)index] = r; image[3 * (size_t)index + 1] = g; image[3 * (size_t)index + 2] = b;
}

// reference :601-691; grad_sigmas / grad_rgbs are pre-zeroed by the caller (raymarching.py:284-285)
__global__ void __launch_bounds__(128) k_composite_train_bwd(
    const float *__restrict__ grad_weights_sum, const float *__restrict__ grad_depth, const float *__restrict__ grad_image,
    const float *__restrict__ sigmas, const float *__restrict__ rgbs, const float *__restrict__ deltas, const int *__restrict__ rays,
    const float *__restrict__ weights_sum, const float *__restrict__ depth, const float *__restrict__ image, uint32_t M, uint32_t N,
    float T_thresh, float *__restrict__ grad_sigmas, float *__restrict__ grad_rgbs) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    const uint32_t index = (uint32_t)rays[3 * (size_t)n], offset = (uint32_t)rays[3 * (size_t)n + 1], num_steps = (uint32_t)rays[3 * (size_t)n + 2];
    if (num_steps == 0 || offset + num_steps > M) return;
    const float gi0 = grad_image[3 * (size_t)index], gi1 = grad_image[3 * (size_t)index + 1], gi2 = grad_image[3 * (size_t)index + 2];
    const float gws = grad_weights_sum[index], gd = grad_depth[index];
    const float r_final = image[3 * (size_t)index], g_final = image[3 * (size_t)index + 1], b_final = image[3 * (size_t)index + 2];
    const float ws_final = weights_sum[index], d_final = depth[index];
    const float *ps = sigmas + offset, *pc = rgbs + 3 * (size_t)offset, *pl = deltas + 2 * (size_t)offset;
    float *gs = grad_sigmas + offset, *gc = grad_rgbs + 3 * (size_t)offset;
    float T = 1.0f, r = 0, g = 0, b = 0, ws = 0, t = 0, d = 0;
    for (uint32_t step = 0; step < num_steps; ++step) {
        const float2 dl = *reinterpret_cast<const float2 *>(pl + 2 * (size_t)step);
        const float c0 = pc[3 * (size_t)step], c1 = pc[3 * (size_t)step + 1], c2 = pc[3 * (size_t)step + 2];
        const float alpha = 1.0f - __expf(-ps[step] * dl.x);
        const float weight = alpha * T;
        r += weight * c0; g += weight * c1; b += weight * c2;
        t += dl.y;
        d += weight * t;
        ws += weight;
        T *= 1.0f - alpha;
        gc[3 * (size_t)step] = gi0 * weight; gc[3 * (size_t)step + 1] = gi1 * weight; gc[3 * (size_t)step + 2] = gi2 * weight;
        gs[step] = dl.x * (gi0 * (T * c0 - (r_final - r)) + gi1 * (T * c1 - (g_final - g)) + gi2 * (T * c2 - (b_final - b)) +
                           gd * (T * t - (d_final - d)) + gws * (1 - ws_final));
        if (T < T_thresh) break;
    }
}

}  // namespace
}  // namespace ngp
