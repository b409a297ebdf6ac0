// nerf_step.cu -- the elementwise glue of one Instant-NGP training / rendering step, fused into a handful
// of kernels so that a step is a fixed sequence of launches on device-resident, fixed-capacity buffers
// (CUDA-graph capturable; no host reads).  This is the "renderer driver" row of the scope table:
//   network:   /root/reference/nerf/network_ff.py:51-74 (sigma = trunc_exp(h[0]), colour-net input
//              [SH16(dir) | geo_feat 15 | 0 pad], rgb = sigmoid), activation.py:5-17 (trunc_exp backward clamp),
//              shencoder/src/shencoder.cu:27-120 (degree-4 SH polynomials)
//   loss:      nerf/utils.py:534 (MSE over rgb, mean), nerf/renderer.py:368 (image + (1 - weights_sum) * bg)
//   optimizer: main_nerf.py:194 (Adam betas (0.9, 0.99) eps 1e-15), :197 (lr * 0.1^(step/30000)),
//              torch.cuda.amp.GradScaler semantics (nerf/utils.py:974-978: unscale, skip on inf/nan, grow/backoff)
// Every kernel here is streaming (HBM/L2 bound): 128-bit accesses, one thread per sample / parameter group.
#include "common.cuh"

namespace ngp {
namespace {

__device__ __forceinline__ uint32_t h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t *>(&h);
}
__device__ __forceinline__ float2 unh2(uint32_t u) { return __half22float2(*reinterpret_cast<const __half2 *>(&u)); }

// x01 = (xyz + bound) / (2 bound)          (gridencoder/grid.py:807)
__global__ void __launch_bounds__(256) k_nerf_x01(const float *__restrict__ xyzs, float bound, float *__restrict__ x01, uint32_t n3, const int *__restrict__ limit) {
    n3 = eff_batch(n3, limit, 3u);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) x01[i] = (xyzs[i] + bound) / (2 * bound);
}

// h [M,16] f16 (sigma net output), dirs [M,3]  ->  sigmas [M] f32 = density_scale * exp(h0),
// cin [M,32] f16 = [SH degree 4 of dir (16) | h1..h15 | 0]
__global__ void __launch_bounds__(256) k_nerf_mid_fwd(const uint4 *__restrict__ h, const float *__restrict__ dirs, float density_scale,
                                                      float *__restrict__ sigmas, uint4 *__restrict__ cin, uint32_t M, const int *__restrict__ limit) {
    M = eff_batch(M, limit);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const uint4 a = __ldg(h + 2 * (size_t)i), b = __ldg(h + 2 * (size_t)i + 1);
    const float2 h01 = unh2(a.x);
    sigmas[i] = density_scale * __expf(h01.x);
    const float x = dirs[3 * (size_t)i], y = dirs[3 * (size_t)i + 1], z = dirs[3 * (size_t)i + 2];
    const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
    uint4 s0, s1;
    s0.x = h2(0.28209479177387814f, -0.48860251190291987f * y);
    s0.y = h2(0.48860251190291987f * z, -0.48860251190291987f * x);
    s0.z = h2(1.0925484305920792f * xy, -1.0925484305920792f * yz);
    s0.w = h2(0.94617469575755997f * z2 - 0.31539156525251999f, -1.0925484305920792f * xz);
    s1.x = h2(0.54627421529603959f * x2 - 0.54627421529603959f * y2, 0.59004358992664352f * y * (-3.0f * x2 + y2));
    s1.y = h2(2.8906114426405538f * xy * z, 0.45704579946446572f * y * (1.0f - 5.0f * z2));
    s1.z = h2(0.3731763325901154f * z * (5.0f * z2 - 3.0f), 0.45704579946446572f * x * (1.0f - 5.0f * z2));
    s1.w = h2(1.4453057213202769f * z * (x2 - y2), 0.59004358992664352f * x * (-x2 + 3.0f * y2));
    // geo_feat = h[1..15] shifted down by one half, then a zero pad
    uint4 g0, g1;
    g0.x = __byte_perm(a.x, a.y, 0x5432); g0.y = __byte_perm(a.y, a.z, 0x5432); g0.z = __byte_perm(a.z, a.w, 0x5432); g0.w = __byte_perm(a.w, b.x, 0x5432);
    g1.x = __byte_perm(b.x, b.y, 0x5432); g1.y = __byte_perm(b.y, b.z, 0x5432); g1.z = __byte_perm(b.z, b.w, 0x5432); g1.w = b.w >> 16;
    uint4 *o = cin + 4 * (size_t)i;
    o[0] = s0; o[1] = s1; o[2] = g0; o[3] = g1;
}

// c [M,16] f16 (colour net output) -> rgbs [M,3] f32 = sigmoid(c[0..2])
__global__ void __launch_bounds__(256) k_nerf_rgb_fwd(const uint2 *__restrict__ c, float *__restrict__ rgbs, uint32_t M, const int *__restrict__ limit) {
    M = eff_batch(M, limit);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const uint2 u = __ldg(c + 4 * (size_t)i);  // first 4 halves of the 32-byte row
    const float2 a = unh2(u.x), b = unh2(u.y);
    // torch.sigmoid of the half tensor under autocast returns half (network_ff.py:73): round like it does
    rgbs[3 * (size_t)i] = __half2float(__float2half_rn(1.0f / (1.0f + __expf(-a.x))));
    rgbs[3 * (size_t)i + 1] = __half2float(__float2half_rn(1.0f / (1.0f + __expf(-a.y))));
    rgbs[3 * (size_t)i + 2] = __half2float(__float2half_rn(1.0f / (1.0f + __expf(-b.x))));
}

// per ray: pred = image + (1 - ws) * bg ; loss += sum_c (pred - gt)^2 / (3 N) ;
// grad_image = scale * 2 (pred - gt) / (3 N) ; grad_ws = -sum_c grad_image_c * bg ; grad_depth = 0
__global__ void __launch_bounds__(256) k_nerf_loss(const float *__restrict__ image, const float *__restrict__ ws, const float *__restrict__ gt,
                                                   float bg, uint32_t N, const float *__restrict__ loss_scale, float *__restrict__ grad_image,
                                                   float *__restrict__ grad_ws, float *__restrict__ grad_depth, float *__restrict__ loss_out) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x;
    float l = 0.f;
    if (n < N) {
        const float s = loss_scale ? *loss_scale : 1.0f;
        const float k = 2.0f / (3.0f * (float)N);
        const float w = 1.0f - ws[n];
        float gw = 0.f;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float d = image[3 * (size_t)n + c] + w * bg - gt[3 * (size_t)n + c];
            l += d * d;
            const float g = s * k * d;
            grad_image[3 * (size_t)n + c] = g;
            gw -= g * bg;
        }
        grad_ws[n] = gw;
        grad_depth[n] = 0.f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    if ((threadIdx.x & 31) == 0 && l != 0.f) atomicAdd(loss_out, l / (3.0f * (float)N));
}

// grad_rgbs [M,3] f32, rgbs [M,3] f32 -> grad_c [M,16] f16: g * rgb (1 - rgb) in columns 0..2, zeros elsewhere
__global__ void __launch_bounds__(256) k_nerf_rgb_bwd(const float *__restrict__ grad_rgbs, const float *__restrict__ rgbs,
                                                      uint4 *__restrict__ grad_c, uint32_t M, const int *__restrict__ limit) {
    M = eff_batch(M, limit);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    float g[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float r = rgbs[3 * (size_t)i + c];
        g[c] = grad_rgbs[3 * (size_t)i + c] * r * (1.0f - r);
    }
    uint4 lo = make_uint4(h2(g[0], g[1]), h2(g[2], 0.f), 0u, 0u);
    grad_c[2 * (size_t)i] = lo;
    grad_c[2 * (size_t)i + 1] = make_uint4(0u, 0u, 0u, 0u);
}

// grad_sigmas [M] f32, h [M,16] f16, grad_cin [M,32] f16 -> grad_h [M,16] f16:
//   col 0 = grad_sigma * density_scale * exp(clamp(h0, -15, 15)) ; cols 1..15 = grad_cin[16..30]
__global__ void __launch_bounds__(256) k_nerf_mid_bwd(const float *__restrict__ grad_sigmas, const uint4 *__restrict__ h,
                                                      const uint4 *__restrict__ grad_cin, float density_scale, uint4 *__restrict__ grad_h,
                                                      uint32_t M, const int *__restrict__ limit) {
    M = eff_batch(M, limit);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= M) return;
    const float h0 = unh2(__ldg(reinterpret_cast<const uint32_t *>(h + 2 * (size_t)i))).x;
    const float g0 = grad_sigmas[i] * density_scale * __expf(fminf(fmaxf(h0, -15.f), 15.f));
    const uint4 a = __ldg(grad_cin + 4 * (size_t)i + 2), b = __ldg(grad_cin + 4 * (size_t)i + 3);  // halves 16..31 = geo grads 0..14, pad
    const __half g0h = __float2half_rn(g0);
    const uint32_t g0u = (uint32_t)(*reinterpret_cast<const unsigned short *>(&g0h));
    uint4 o0, o1;  // shift up by one half, insert g0 at column 0
    o0.x = g0u | (a.x << 16); o0.y = __byte_perm(a.x, a.y, 0x5432); o0.z = __byte_perm(a.y, a.z, 0x5432); o0.w = __byte_perm(a.z, a.w, 0x5432);
    o1.x = __byte_perm(a.w, b.x, 0x5432); o1.y = __byte_perm(b.x, b.y, 0x5432); o1.z = __byte_perm(b.y, b.z, 0x5432); o1.w = __byte_perm(b.z, b.w, 0x5432);
    grad_h[2 * (size_t)i] = o0;
    grad_h[2 * (size_t)i + 1] = o1;
}



// SDF regression (sdf/utils.py:217-225 with loss.py:7-15 mape_loss, mean): pred = h[:, 0] (fp16 MLP output row of 16),
//   loss += |pred - y| / (|y| + 1e-2) / B ;  grad_h[:, 0] = scale * sign(pred - y) / (|y| + 1e-2) / B, other columns 0
__global__ void __launch_bounds__(256) k_sdf_loss(const uint4 *__restrict__ h, const float *__restrict__ target, uint32_t B,
                                                  const float *__restrict__ loss_scale, uint4 *__restrict__ grad_h, float *__restrict__ loss_out,
                                                  const int *__restrict__ limit) {
    B = eff_batch(B, limit);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float l = 0.f;
    if (i < B) {
        const float pred = unh2(__ldg(reinterpret_cast<const uint32_t *>(h + 2 * (size_t)i))).x;
        const float y = target[i];
        const float w = 1.0f / (fabsf(y) + 1e-2f), d = pred - y;
        l = fabsf(d) * w;
        const float s = loss_scale ? *loss_scale : 1.0f;
        const float g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
        grad_h[2 * (size_t)i] = make_uint4(h2(s * g * w / (float)B, 0.f), 0u, 0u, 0u);
        grad_h[2 * (size_t)i + 1] = make_uint4(0u, 0u, 0u, 0u);
    }
    // one atomic per block (2^20 points would otherwise queue 32768 same-address atomics)
    __shared__ float s_part[8];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = l;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s_part[w];
        if (t != 0.f) atomicAdd(loss_out, t / (float)B);
    }
}

// ---- optimizer ---------------------------------------------------------------------------------------------
// Device-resident step state (float[16], ints bit-cast), so a captured step needs no host-side scalars:
//   [0] optimizer step count (int)   [1] lr for this step      [2] 1 - beta1^t        [3] 1 - beta2^t
//   [4] loss scale                   [5] growth tracker (int)  [6] found_inf (int)    [7] loss accumulator
//   [8] loss of the last finished step                         [9] lr0                [10] decay steps  [11] decay factor
//   [12] growth interval (int)
enum { ST_STEP = 0, ST_LR, ST_BC1, ST_BC2, ST_SCALE, ST_GROWTH, ST_FOUND_INF, ST_LOSS, ST_LAST_LOSS, ST_LR0, ST_DECAY_STEPS, ST_DECAY, ST_GROWTH_INTERVAL, ST_ITER };

// start of a step: clear the loss accumulator and the inf flag; derive lr / bias corrections for step t = count + 1
__global__ void k_step_begin(float *__restrict__ st, float beta1, float beta2) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int t = __float_as_int(st[ST_STEP]) + 1;
    st[ST_LOSS] = 0.f;
    st[ST_FOUND_INF] = __int_as_float(0);
    // LambdaLR(optimizer, lambda iter: 0.1 ** min(iter / 30000, 1)) (main_nerf.py:197): the scheduler steps once per ITERATION,
    // also when GradScaler skipped the optimizer step (nerf/utils.py:974-981) -- ST_ITER counts iterations, ST_STEP the optimizer
    // steps Adam's bias correction sees (torch.optim.Adam's own counter does not advance on a skipped step)
    const float frac = fminf((float)__float_as_int(st[ST_ITER]) / st[ST_DECAY_STEPS], 1.0f);
    st[ST_LR] = st[ST_LR0] * powf(st[ST_DECAY], frac);
    st[ST_BC1] = 1.0f - powf(beta1, (float)t);
    st[ST_BC2] = 1.0f - powf(beta2, (float)t);
}
// end of a step: GradScaler.update() + step count
__global__ void k_step_end(float *__restrict__ st) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int found = __float_as_int(st[ST_FOUND_INF]);
    int growth = __float_as_int(st[ST_GROWTH]);
    if (found) {
        st[ST_SCALE] *= 0.5f;
        growth = 0;
    } else {
        st[ST_STEP] = __int_as_float(__float_as_int(st[ST_STEP]) + 1);
        if (++growth >= __float_as_int(st[ST_GROWTH_INTERVAL])) {
            st[ST_SCALE] *= 2.0f;
            growth = 0;
        }
    }
    st[ST_GROWTH] = __int_as_float(growth);
    st[ST_ITER] = __int_as_float(__float_as_int(st[ST_ITER]) + 1);
    st[ST_LAST_LOSS] = st[ST_LOSS];
}

// any non-finite gradient -> found_inf (GradScaler.unscale_'s check).  g16: n16 halves (multiple of 8), g32: n32 floats (multiple of 4)
__global__ void __launch_bounds__(256) k_found_inf(const uint4 *__restrict__ g16, size_t n16_8, const float4 *__restrict__ g32, size_t n32_4,
                                                   float *__restrict__ st) {
    bool bad = false;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16_8; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 u = __ldg(g16 + i);
        // fp16 inf/nan: exponent bits all ones
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) bad |= ((w[k] & 0x7C00u) == 0x7C00u) | ((w[k] & 0x7C000000u) == 0x7C000000u);
    }
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n32_4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = __ldg(g32 + i);
        bad |= !(isfinite(v.x) && isfinite(v.y) && isfinite(v.z) && isfinite(v.w));
    }
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) st[ST_FOUND_INF] = __int_as_float(1);
}

// Adam over one flat fp32 parameter vector p[0 .. n16 + n32): the first n16 parameters (hash table) have fp16
// gradients and an fp16 shadow copy (what the encoder reads under AMP, gridencoder/grid.py:55-56 without the
// per-forward 73 MB cast), the last n32 (MLP weights) have fp32 gradients.  Gradients carry the loss scale;
// they are consumed AND cleared here (ready for the next backward's accumulation).  torch.optim.Adam formula:
//   m = b1 m + (1-b1) g ; v = b2 v + (1-b2) g^2 ; p -= lr / bc1 * m / (sqrt(v) / sqrt(bc2) + eps)
__global__ void __launch_bounds__(256) k_adam(float4 *__restrict__ p, float4 *__restrict__ m, float4 *__restrict__ v, uint2 *__restrict__ g16,
                                              float4 *__restrict__ g32, uint2 *__restrict__ shadow, size_t n16_4, size_t n32_4,
                                              const float *__restrict__ st, float beta1, float beta2, float eps, float grad_div, float wd32) {
    const bool skip = __float_as_int(st[ST_FOUND_INF]) != 0;
    const float inv = 1.0f / (st[ST_SCALE] * grad_div);
    const float lr = st[ST_LR], bc1 = st[ST_BC1], rbc2 = rsqrtf(st[ST_BC2]);
    // grid-stride: the launch may be capped to a few blocks per SM so that the kernel shares the SMs with the ray marcher
    // of the next step (NGPTrainer overlap_optimizer); it is HBM-bound and saturates the memory system well below full occupancy
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16_4 + n32_4; i += (size_t)gridDim.x * blockDim.x) {
        float g[4];
        if (i < n16_4) {
            const uint2 u = g16[i];
            const float2 a = unh2(u.x), b = unh2(u.y);
            g[0] = a.x; g[1] = a.y; g[2] = b.x; g[3] = b.y;
            if (u.x | u.y) g16[i] = make_uint2(0u, 0u);
        } else {
            const float4 t = g32[i - n16_4];
            g[0] = t.x; g[1] = t.y; g[2] = t.z; g[3] = t.w;
            g32[i - n16_4] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (skip) continue;
        float4 pp = p[i], mm = m[i], vv = v[i];
        float *pf = &pp.x, *mf = &mm.x, *vf = &vv.x;
        const float wd = i < n16_4 ? 0.f : wd32;  // torch.optim.Adam weight_decay (L2 term added to the gradient), MLP group only
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float gk = g[k] * inv + wd * pf[k];
            mf[k] = beta1 * mf[k] + (1.0f - beta1) * gk;
            vf[k] = beta2 * vf[k] + (1.0f - beta2) * gk * gk;
            pf[k] -= (lr / bc1) * mf[k] / (sqrtf(vf[k]) * rbc2 + eps);
        }
        p[i] = pp; m[i] = mm; v[i] = vv;
        if (i < n16_4) shadow[i] = make_uint2(h2(pp.x, pp.y), h2(pp.z, pp.w));
    }
}

// fp32 -> fp16 shadow (initialisation / checkpoint load)
__global__ void __launch_bounds__(256) k_cast_f16(const float4 *__restrict__ src, uint2 *__restrict__ dst, size_t n4) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n4) return;
    const float4 v = __ldg(src + i);
    dst[i] = make_uint2(h2(v.x, v.y), h2(v.z, v.w));
}

// ---- ray generation (nerf/utils.py:52-136 get_rays, the part after the pixel indices are chosen) ---------------------
// One thread per ray of pose blockIdx.y: pixel centre -> camera direction ((i - cx)/fx, (j - cy)/fy, 1) normalised ->
// world direction (directions @ R^T), origin = camera position.  inds: pixel indices [N] shared by all poses
// (inds_stride = 0), [B, N] (inds_stride = N), or NULL = every pixel in row-major order.
__global__ void __launch_bounds__(256) k_get_rays(const float *__restrict__ poses, float fx, float fy, float cx, float cy, uint32_t W,
                                                  const int *__restrict__ inds, uint32_t inds_stride, uint32_t N, float *__restrict__ rays_o,
                                                  float *__restrict__ rays_d) {
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x, b = blockIdx.y;
    if (n >= N) return;
    const float *P = poses + (size_t)b * 16;
    const uint32_t pix = inds ? (uint32_t)inds[(size_t)b * inds_stride + n] : n;
    const float i = (float)(pix % W) + 0.5f, j = (float)(pix / W) + 0.5f;
    const float xs = __fdiv_rn(__fsub_rn(i, cx), fx), ys = __fdiv_rn(__fsub_rn(j, cy), fy);
    const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(xs, xs), __fmul_rn(ys, ys)), 1.0f));
    const float dx = __fdiv_rn(xs, nrm), dy = __fdiv_rn(ys, nrm), dz = __fdiv_rn(1.0f, nrm);
    const size_t o = ((size_t)b * N + n) * 3;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        rays_d[o + k] = __fadd_rn(__fadd_rn(__fmul_rn(dx, P[4 * k]), __fmul_rn(dy, P[4 * k + 1])), __fmul_rn(dz, P[4 * k + 2]));
        rays_o[o + k] = P[4 * k + 3];
    }
}

// ---- density grid maintenance (nerf/renderer.py:498-598) ----------------------------------------------------
// grid = max(grid * decay, tmp) where both >= 0 ; sum of clamp(grid, 0) -> mean density.
// The sum is carried in DOUBLE (per thread, per warp, and in the one global accumulator): the reference takes
// torch.mean of 2 M+ cells (renderer.py:584) and thresholds the grid with it (:589), so the mean has to be the
// correctly rounded fp32 mean whatever the order of the additions -- a chain of fp32 atomics onto a ~1e7 running sum
// was off by 2.5e-6 relative (measured), enough to flip cells that sit next to the threshold.
__global__ void __launch_bounds__(256) k_density_ema(float *__restrict__ grid, const float *__restrict__ tmp, float decay, uint32_t n,
                                                     double *__restrict__ sum_out) {
    double s = 0.0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float g = grid[i];
        const float t = tmp[i];
        if (g >= 0.f && t >= 0.f) {
            g = fmaxf(g * decay, t);
            grid[i] = g;
        }
        s += (double)fmaxf(g, 0.f);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    __shared__ double s_part[8];
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s_part[w];
        atomicAdd(sum_out, t);
    }
}
// packbits with a device-resident threshold: thresh = min(sum / n, density_thresh)  (renderer.py:584-590)
__global__ void __launch_bounds__(256) k_packbits_mean(const float4 *__restrict__ grid4, uint32_t N4, const double *__restrict__ sum, double inv_n,
                                                       float density_thresh, uint32_t *__restrict__ bitfield32) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N4) return;
    const float thresh = fminf((float)(*sum * inv_n), density_thresh);  // the fp32 mean, rounded once
    uint32_t bits = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float4 v = __ldg(grid4 + 8 * (size_t)n + i);
        bits |= (v.x > thresh ? 1u : 0u) << (4 * i) | (v.y > thresh ? 1u : 0u) << (4 * i + 1) | (v.z > thresh ? 1u : 0u) << (4 * i + 2) |
                (v.w > thresh ? 1u : 0u) << (4 * i + 3);
    }
    bitfield32[n] = bits;
}

// ---- spatial binning of an UNORDERED batch (the SDF step; sdf/provider.py hands out surface / uniform samples in random order) ----
// With random order every (point, level) is its own L2 reduction in the table-gradient kernel, and on the coarse levels
// thousands of them meet in the same entries: the L2 atomic unit serialises per address (measured on 2^20 torus samples:
// 59-92 us per coarse level against 29 us for a fine one, half of the whole backward in levels 0-3).  The loss is a mean
// over the batch, so the order of its points is free: one counting sort by the Morton code of a 2^bits cube grid puts the
// points of a cell next to each other, and the segmented warp reduction of k_grid_bwd_v3 -- written for samples marched
// along a ray -- then merges them before they reach L2; the gathers of the forward pass become L1 hits the same way.
__device__ __forceinline__ uint32_t spread3_10(uint32_t v) {  // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}

__global__ void __launch_bounds__(256) k_bin_hist(const float *__restrict__ points, uint32_t B, float bound, uint32_t bits, uint32_t *__restrict__ hist,
                                                  uint32_t *__restrict__ key, uint32_t *__restrict__ rank) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const float R = (float)(1u << bits), inv = 0.5f / bound;
    uint32_t c[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float u = fminf(fmaxf((points[3 * (size_t)b + d] + bound) * inv, 0.f), 1.f);  // NaN -> 0
        c[d] = min((1u << bits) - 1u, (uint32_t)(u * R));
    }
    const uint32_t k = spread3_10(c[0]) | (spread3_10(c[1]) << 1) | (spread3_10(c[2]) << 2);
    key[b] = k;
    rank[b] = atomicAdd(hist + k, 1u);  // my slot inside the bin
}

// 8192 bins per block: sum them
__global__ void __launch_bounds__(1024) k_bin_blocksum(const uint32_t *__restrict__ hist, uint32_t *__restrict__ partial) {
    __shared__ uint32_t s_w[32];
    const uint4 *h4 = reinterpret_cast<const uint4 *>(hist) + ((size_t)blockIdx.x * 1024 + threadIdx.x) * 2;
    const uint4 a = h4[0], b = h4[1];
    uint32_t v = a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
    v = __reduce_add_sync(0xffffffffu, v);
    if ((threadIdx.x & 31u) == 0u) s_w[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        v = __reduce_add_sync(0xffffffffu, s_w[threadIdx.x]);
        if (threadIdx.x == 0) partial[blockIdx.x] = v;
    }
}

// exclusive prefix sum of the histogram -> start[]; the histogram is left zeroed for the next call
__global__ void __launch_bounds__(1024) k_bin_scan(uint32_t *__restrict__ hist, const uint32_t *__restrict__ partial, uint32_t *__restrict__ start) {
    __shared__ uint32_t s_w[32];
    __shared__ uint32_t s_base;
    uint32_t before = 0;
    for (uint32_t i = threadIdx.x; i < blockIdx.x; i += blockDim.x) before += partial[i];
    before = __reduce_add_sync(0xffffffffu, before);
    if ((threadIdx.x & 31u) == 0u) s_w[threadIdx.x >> 5] = before;
    __syncthreads();
    if (threadIdx.x < 32) {
        const uint32_t t = __reduce_add_sync(0xffffffffu, s_w[threadIdx.x]);
        if (threadIdx.x == 0) s_base = t;
    }
    __syncthreads();
    const uint32_t base = s_base;
    uint4 *h4 = reinterpret_cast<uint4 *>(hist) + ((size_t)blockIdx.x * 1024 + threadIdx.x) * 2;
    const uint4 a = h4[0], b = h4[1];
    h4[0] = make_uint4(0u, 0u, 0u, 0u);
    h4[1] = make_uint4(0u, 0u, 0u, 0u);
    const uint32_t mine = a.x + a.y + a.z + a.w + b.x + b.y + b.z + b.w;
    uint32_t inc = mine;  // inclusive scan over the warp
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, inc, d);
        if ((threadIdx.x & 31u) >= (uint32_t)d) inc += u;
    }
    __syncthreads();  // s_w is reused
    if ((threadIdx.x & 31u) == 31u) s_w[threadIdx.x >> 5] = inc;
    __syncthreads();
    uint32_t wbase = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) wbase += s_w[w];
    uint32_t e = base + wbase + inc - mine;
    uint4 o0, o1;
    o0.x = e; e += a.x; o0.y = e; e += a.y; o0.z = e; e += a.z; o0.w = e; e += a.w;
    o1.x = e; e += b.x; o1.y = e; e += b.y; o1.z = e; e += b.z; o1.w = e;
    uint4 *s4 = reinterpret_cast<uint4 *>(start) + ((size_t)blockIdx.x * 1024 + threadIdx.x) * 2;
    s4[0] = o0;
    s4[1] = o1;
}

__global__ void __launch_bounds__(256) k_bin_scatter(const float *__restrict__ points, const float *__restrict__ values, uint32_t B,
                                                     const uint32_t *__restrict__ key, const uint32_t *__restrict__ rank,
                                                     const uint32_t *__restrict__ start, float *__restrict__ points_out, float *__restrict__ values_out) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint32_t dst = start[key[b]] + rank[b];
    points_out[3 * (size_t)dst] = points[3 * (size_t)b];
    points_out[3 * (size_t)dst + 1] = points[3 * (size_t)b + 1];
    points_out[3 * (size_t)dst + 2] = points[3 * (size_t)b + 2];
    if (values) values_out[dst] = values[b];
}

}  // namespace
}  // namespace ngp

extern "C" {

#define NGP_GRID1(n) ngp::div_round_up((uint32_t)(n), 256u), 256, 0, st

int ngp_nerf_x01(const float *xyzs, float bound, float *x01, uint32_t M, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (M == 0) return NGP_OK;
    NGP_CHECK_ARG(xyzs && x01, "nerf_x01: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_nerf_x01", st); k_nerf_x01<<<NGP_GRID1(3 * M)>>>(xyzs, bound, x01, 3 * M, batch_limit()); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_nerf_mid_forward(const void *h, const float *dirs, float density_scale, float *sigmas, void *cin, uint32_t M, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (M == 0) return NGP_OK;
    NGP_CHECK_ARG(h && dirs && sigmas && cin, "nerf_mid_forward: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_nerf_mid_fwd", st); k_nerf_mid_fwd<<<NGP_GRID1(M)>>>((const uint4 *)h, dirs, density_scale, sigmas, (uint4 *)cin, M, batch_limit()); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_nerf_rgb_forward(const void *c, float *rgbs, uint32_t M, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (M == 0) return NGP_OK;
    NGP_CHECK_ARG(c && rgbs, "nerf_rgb_forward: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_nerf_rgb_fwd", st); k_nerf_rgb_fwd<<<NGP_GRID1(M)>>>((const uint2 *)c, rgbs, M, batch_limit()); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_nerf_loss(const float *image, const float *weights_sum, const float *gt, float bg, uint32_t N, const float *loss_scale,
                  float *grad_image, float *grad_weights_sum, float *grad_depth, float *loss_out, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(image && weights_sum && gt && grad_image && grad_weights_sum && grad_depth && loss_out, "nerf_loss: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_nerf_loss", st); k_nerf_loss<<<NGP_GRID1(N)>>>(image, weights_sum, gt, bg, N, loss_scale, grad_image, grad_weights_sum, grad_depth, loss_out); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_nerf_rgb_backward(const float *grad_rgbs, const float *rgbs, void *grad_c, uint32_t M, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (M == 0) return NGP_OK;
    NGP_CHECK_ARG(grad_rgbs && rgbs && grad_c, "nerf_rgb_backward: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_nerf_rgb_bwd", st); k_nerf_rgb_bwd<<<NGP_GRID1(M)>>>(grad_rgbs, rgbs, (uint4 *)grad_c, M, batch_limit()); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_nerf_mid_backward(const float *grad_sigmas, const void *h, const void *grad_cin, float density_scale, void *grad_h, uint32_t M,
                          void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (M == 0) return NGP_OK;
    NGP_CHECK_ARG(grad_sigmas && h && grad_cin && grad_h, "nerf_mid_backward: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_nerf_mid_bwd", st); k_nerf_mid_bwd<<<NGP_GRID1(M)>>>(grad_sigmas, (const uint4 *)h, (const uint4 *)grad_cin, density_scale, (uint4 *)grad_h, M, batch_limit()); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_sdf_loss(const void *h, const float *target, uint32_t B, const float *loss_scale, void *grad_h, float *loss_out, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (B == 0) return NGP_OK;
    NGP_CHECK_ARG(h && target && grad_h && loss_out, "sdf_loss: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_sdf_loss", st); k_sdf_loss<<<NGP_GRID1(B)>>>((const uint4 *)h, target, B, loss_scale, (uint4 *)grad_h, loss_out, batch_limit()); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_bin_points(const float *points, const float *values, uint32_t B, float bound, uint32_t bits, void *workspace, size_t workspace_bytes,
                   float *points_out, float *values_out, void *stream) {
    using namespace ngp;
    if (B == 0) return NGP_OK;
    NGP_CHECK_ARG(points && points_out && workspace && (values == nullptr) == (values_out == nullptr), "bin_points: null pointer argument");
    NGP_CHECK_ARG(bits >= 5 && bits <= 7 && bound > 0.f, "bin_points: bits must be 5, 6 or 7 (got %u) and bound > 0", bits);
    NGP_CHECK_ARG(points != points_out, "bin_points: the copy is not in place");
    const size_t nbins = (size_t)1 << (3 * bits), nblk = nbins / 8192;
    NGP_CHECK_ARG(workspace_bytes >= ngp_bin_points_workspace(B, bits) && ((uintptr_t)workspace & 15) == 0,
                  "bin_points: workspace needs %zu bytes, 16-byte aligned", ngp_bin_points_workspace(B, bits));
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t *hist = (uint32_t *)workspace, *start = hist + nbins, *partial = start + nbins, *key = partial + ((nblk + 3) & ~(size_t)3), *rank = key + B;
    { NGP_PROF("k_bin_hist", st); k_bin_hist<<<NGP_GRID1(B)>>>(points, B, bound, bits, hist, key, rank); }
    { NGP_PROF("k_bin_blocksum", st); k_bin_blocksum<<<(uint32_t)nblk, 1024, 0, st>>>(hist, partial); }
    { NGP_PROF("k_bin_scan", st); k_bin_scan<<<(uint32_t)nblk, 1024, 0, st>>>(hist, partial, start); }
    { NGP_PROF("k_bin_scatter", st); k_bin_scatter<<<NGP_GRID1(B)>>>(points, values, B, key, rank, start, points_out, values_out); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

size_t ngp_bin_points_workspace(uint32_t B, uint32_t bits) {
    const size_t nbins = (size_t)1 << (3 * bits), nblk = nbins / 8192;
    return (2 * nbins + ((nblk + 3) & ~(size_t)3) + 2 * (size_t)B) * sizeof(uint32_t);
}

int ngp_step_begin(float *state, float beta1, float beta2, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(state, "step_begin: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_step_begin", st); k_step_begin<<<1, 32, 0, st>>>(state, beta1, beta2); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_step_end(float *state, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(state, "step_end: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_step_end", st); k_step_end<<<1, 32, 0, st>>>(state); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_found_inf(const void *g16, size_t n16, const float *g32, size_t n32, float *state, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(state && (n16 == 0 || g16) && (n32 == 0 || g32), "found_inf: null pointer argument");
    NGP_CHECK_ARG(n16 % 8 == 0 && n32 % 4 == 0, "found_inf: n16 must be a multiple of 8 and n32 of 4");
    if (n16 + n32 == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t work = n16 / 8 + n32 / 4;
    uint32_t grid = (uint32_t)((work + 255) / 256);
    const uint32_t cap = (uint32_t)sm_count() * 16;
    if (grid > cap) grid = cap;
    { NGP_PROF("k_found_inf", st); k_found_inf<<<grid, 256, 0, st>>>((const uint4 *)g16, n16 / 8, (const float4 *)g32, n32 / 4, state); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_adam_step(float *params, float *exp_avg, float *exp_avg_sq, void *g16, float *g32, void *shadow16, size_t n16, size_t n32,
                  const float *state, float beta1, float beta2, float eps, float grad_div, float weight_decay32, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    NGP_CHECK_ARG(params && exp_avg && exp_avg_sq && state && (n16 == 0 || (g16 && shadow16)) && (n32 == 0 || g32), "adam_step: null pointer argument");
    NGP_CHECK_ARG(n16 % 4 == 0 && n32 % 4 == 0, "adam_step: segment sizes must be multiples of 4");
    if (n16 + n32 == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t n4 = (n16 + n32) / 4;
    uint32_t grid = (uint32_t)((n4 + 255) / 256);
    const int per_sm = grid_cap_blocks_per_sm();  // ngp_call_opts: share the SMs with another stream's kernels
    if (per_sm > 0 && grid > (uint32_t)(per_sm * sm_count())) grid = (uint32_t)(per_sm * sm_count());
    { NGP_PROF("k_adam", st); k_adam<<<grid, 256, 0, st>>>((float4 *)params, (float4 *)exp_avg, (float4 *)exp_avg_sq, (uint2 *)g16, (float4 *)g32,
                                                           (uint2 *)shadow16, n16 / 4, n32 / 4, state, beta1, beta2, eps, grad_div, weight_decay32); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_cast_f16(const float *src, void *dst, size_t n, void *stream) {
    using namespace ngp;
    if (n == 0) return NGP_OK;
    NGP_CHECK_ARG(src && dst && n % 4 == 0, "cast_f16: null pointer argument or n not a multiple of 4");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_cast_f16", st); k_cast_f16<<<(uint32_t)((n / 4 + 255) / 256), 256, 0, st>>>((const float4 *)src, (uint2 *)dst, n / 4); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_get_rays(const float *poses, uint32_t B, float fx, float fy, float cx, float cy, uint32_t W, const int *inds, uint32_t inds_stride,
                 uint32_t N, float *rays_o, float *rays_d, void *stream) {
    using namespace ngp;
    if (B == 0 || N == 0) return NGP_OK;
    NGP_CHECK_ARG(poses && rays_o && rays_d, "get_rays: null pointer argument");
    NGP_CHECK_ARG(W > 0 && fx != 0.f && fy != 0.f && B <= 65535, "get_rays: bad image width / focal length / pose count");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_get_rays", st); k_get_rays<<<dim3(div_round_up(N, 256u), B, 1), 256, 0, st>>>(poses, fx, fy, cx, cy, W, inds, inds_stride, N, rays_o, rays_d); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_density_ema(float *grid, const float *tmp, float decay, uint32_t n, double *sum_out, void *stream) {
    using namespace ngp;
    if (n == 0) return NGP_OK;
    NGP_CHECK_ARG(grid && tmp && sum_out, "density_ema: null pointer argument");
    cudaStream_t st = (cudaStream_t)stream;
    uint32_t g = div_round_up(n, 256u);
    const uint32_t cap = (uint32_t)sm_count() * 16;
    if (g > cap) g = cap;
    NGP_CHECK_CUDA(cudaMemsetAsync(sum_out, 0, sizeof(double), st));
    { NGP_PROF("k_density_ema", st); k_density_ema<<<g, 256, 0, st>>>(grid, tmp, decay, n, sum_out); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_packbits_mean(const float *grid, uint32_t N, const double *sum, double inv_count, float density_thresh, uint8_t *bitfield, void *stream) {
    using namespace ngp;
    if (N == 0) return NGP_OK;
    NGP_CHECK_ARG(grid && sum && bitfield && N % 4 == 0, "packbits_mean: null pointer argument or N not a multiple of 4");
    cudaStream_t st = (cudaStream_t)stream;
    { NGP_PROF("k_packbits_mean", st); k_packbits_mean<<<NGP_GRID1(N / 4)>>>((const float4 *)grid, N / 4, sum, inv_count, density_thresh, (uint32_t *)bitfield); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

}  // extern "C"
