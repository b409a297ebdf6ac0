// raymarch.cu -- occupancy-bitfield ray marcher, compositors and helpers for sm_100a.
//
// Contract: /root/reference/raymarching/src/raymarching.cu (utils :91-289, training :311-702,
// inference :709-923) behind the C ABI of include/ngp_b200.h.
//
// Bit-exactness.  Sample positions, deltas, counts and bitfields must equal the reference's.
// Every float operation of the stepping rule is therefore spelled with explicit rounding
// intrinsics (__fmaf_rn / __fmul_rn / __fadd_rn / __fdiv_rn) in exactly the shape nvcc's default
// -fmad=true contraction gives the reference source, so the result does not depend on this
// file's compile flags.  The reference's FP64 islands (`0.5 * (x*rb + 1) * H`, `dt * H * 0.5`)
// are exact products followed by one rounding to float; they are reproduced in fp32 as
// RN(f * H) * 0.5f, which is identical for H < 2^24 (no DMUL / F2F on the hot loop).
//
// Placement.  The reference reserves output segments with atomicAdd in arrival order
// (:405-406).  Here: pass 1 counts per ray and block-scans, the last block to finish scans the
// block totals (threadfence pattern), pass 2 writes -- so rays[n] = (n, offset_n, count_n) with
// offsets in ray order, deterministic, and each block's samples form one contiguous range.
#include <float.h>

#include "common.cuh"

namespace ngp {
namespace {

constexpr float kSqrt3 = 1.7320508075688772f;
constexpr float kRPi = 0.3183098861837907f;
constexpr int kMarchBlock = 128;

__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }
__device__ __forceinline__ float signf(float x) { return copysignf(1.0f, x); }

__host__ __device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__host__ __device__ __forceinline__ uint32_t morton3D_1(uint32_t x, uint32_t y, uint32_t z) {
    return expand_bits(x) | (expand_bits(y) << 1) | (expand_bits(z) << 2);
}
__host__ __device__ __forceinline__ uint32_t compact_bits(uint32_t x) {
    x = x & 0x49249249;
    x = (x | (x >> 2)) & 0xc30c30c3;
    x = (x | (x >> 4)) & 0x0f00f00f;
    x = (x | (x >> 8)) & 0xff0000ff;
    x = (x | (x >> 16)) & 0x0000ffff;
    return x;
}

// exponent of frexpf clamped to [0, max_cascade-1] (reference :42-54)
__device__ __forceinline__ int mip_of(float mx, float max_cascade) {
    int e;
    frexpf(mx, &e);
    return (int)fminf(max_cascade - 1, fmaxf(0.0f, (float)e));
}

struct ray_t {
    float ox, oy, oz, dx, dy, dz, rdx, rdy, rdz;
};
struct march_consts_t {
    float bound, dt_gamma, dt_min, dt_max, rH, H3, Hf, Cf, Hm1;
    const uint8_t *grid;
};

__device__ __forceinline__ march_consts_t make_consts(float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H,
                                                      const uint8_t *grid) {
    march_consts_t c;
    c.bound = bound;
    c.dt_gamma = dt_gamma;
    c.dt_min = __fdiv_rn(2 * kSqrt3, (float)max_steps);
    c.dt_max = __fdiv_rn(__fmul_rn(2 * kSqrt3, (float)(1 << (C - 1))), (float)H);
    c.rH = __fdiv_rn(1.0f, (float)H);
    c.H3 = (float)(H * H * H);
    c.Hf = (float)H;
    c.Cf = (float)C;
    c.Hm1 = (float)(H - 1);
    c.grid = grid;
    return c;
}
__device__ __forceinline__ ray_t load_ray(const float *__restrict__ rays_o, const float *__restrict__ rays_d, uint32_t n) {
    ray_t r;
    r.ox = rays_o[3 * (size_t)n]; r.oy = rays_o[3 * (size_t)n + 1]; r.oz = rays_o[3 * (size_t)n + 2];
    r.dx = rays_d[3 * (size_t)n]; r.dy = rays_d[3 * (size_t)n + 1]; r.dz = rays_d[3 * (size_t)n + 2];
    r.rdx = __fdiv_rn(1.0f, r.dx); r.rdy = __fdiv_rn(1.0f, r.dy); r.rdz = __fdiv_rn(1.0f, r.dz);
    return r;
}

__device__ __forceinline__ int cell_of(float p, float mip_rbound, const march_consts_t &c) {
    // (int)clamp(0.5 * (p*rb + 1) * H, 0, H-1) with the double island folded (see file header)
    const float v = __fmul_rn(0.5f, __fmul_rn(__fmaf_rn(p, mip_rbound, 1.0f), c.Hf));
    return (int)clampf(v, 0.0f, c.Hm1);
}
__device__ __forceinline__ float exit_t(int n, float d, float rd, float p, float mip_bound, const march_consts_t &c) {
    // (((n + 0.5 + 0.5*sign(d)) * rH * 2 - 1) * mip_bound - p) * rd     (reference :390-392)
    const float a = __fadd_rn(__fadd_rn((float)n, 0.5f), __fmul_rn(0.5f, signf(d)));
    const float q = __fmaf_rn(__fmul_rn(a, c.rH), 2.0f, -1.0f);
    return __fmul_rn(__fmaf_rn(q, mip_bound, -p), rd);
}

// One iteration of the marching loop (reference :360-399).  Returns true when the cell at t is
// occupied (x,y,z,dt valid, t untouched); otherwise advances t beyond the empty voxel.
__device__ __forceinline__ bool probe(const ray_t &r, const march_consts_t &c, float &t, float &x, float &y, float &z, float &dt) {
    x = clampf(__fmaf_rn(t, r.dx, r.ox), -c.bound, c.bound);
    y = clampf(__fmaf_rn(t, r.dy, r.oy), -c.bound, c.bound);
    z = clampf(__fmaf_rn(t, r.dz, r.oz), -c.bound, c.bound);
    dt = clampf(__fmul_rn(t, c.dt_gamma), c.dt_min, c.dt_max);

    const int level = max(mip_of(fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z))), c.Cf),
                          mip_of(__fmul_rn(__fmul_rn(dt, c.Hf), 0.5f), c.Cf));
    const float mip_bound = fminf(scalbnf(1.0f, level), c.bound);
    const float mip_rbound = __fdiv_rn(1.0f, mip_bound);

    const int nx = cell_of(x, mip_rbound, c), ny = cell_of(y, mip_rbound, c), nz = cell_of(z, mip_rbound, c);
    // the reference forms the bit index in FLOAT: level * H3 + morton (:339,378)
    const uint32_t index = (uint32_t)__fmaf_rn((float)level, c.H3, (float)morton3D_1((uint32_t)nx, (uint32_t)ny, (uint32_t)nz));
    const bool occ = __ldg(c.grid + (index >> 3)) & (1u << (index & 7u));
    if (occ) return true;

    const float tx = exit_t(nx, r.dx, r.rdx, x, mip_bound, c);
    const float ty = exit_t(ny, r.dy, r.rdy, y, mip_bound, c);
    const float tz = exit_t(nz, r.dz, r.rdz, z, mip_bound, c);
    const float tt = __fadd_rn(t, fmaxf(0.0f, fminf(tx, fminf(ty, tz))));
    do {
        t = __fadd_rn(t, clampf(__fmul_rn(t, c.dt_gamma), c.dt_min, c.dt_max));
    } while (t < tt);
    return false;
}

// ------------------------------------------------------------------------------------------
// utils
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_near_far(const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                                                  const float *__restrict__ aabb, uint32_t N, float min_near,
                                                  float *__restrict__ nears, float *__restrict__ fars) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    const ray_t r = load_ray(rays_o, rays_d, n);
    float near = __fmul_rn(__fsub_rn(aabb[0], r.ox), r.rdx), far = __fmul_rn(__fsub_rn(aabb[3], r.ox), r.rdx), tmp;
    if (near > far) { tmp = near; near = far; far = tmp; }
    float ny = __fmul_rn(__fsub_rn(aabb[1], r.oy), r.rdy), fy = __fmul_rn(__fsub_rn(aabb[4], r.oy), r.rdy);
    if (ny > fy) { tmp = ny; ny = fy; fy = tmp; }
    if (near > fy || ny > far) { nears[n] = fars[n] = FLT_MAX; return; }
    if (ny > near) near = ny;
    if (fy < far) far = fy;
    float nz = __fmul_rn(__fsub_rn(aabb[2], r.oz), r.rdz), fz = __fmul_rn(__fsub_rn(aabb[5], r.oz), r.rdz);
    if (nz > fz) { tmp = nz; nz = fz; fz = tmp; }
    if (near > fz || nz > far) { nears[n] = fars[n] = FLT_MAX; return; }
    if (nz > near) near = nz;
    if (fz < far) far = fz;
    if (near < min_near) near = min_near;
    nears[n] = near;
    fars[n] = far;
}

__global__ void __launch_bounds__(256) k_sph_from_ray(const float *__restrict__ rays_o, const float *__restrict__ rays_d,
                                                      float radius, uint32_t N, float *__restrict__ coords) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    const float ox = rays_o[3 * (size_t)n], oy = rays_o[3 * (size_t)n + 1], oz = rays_o[3 * (size_t)n + 2];
    const float dx = rays_d[3 * (size_t)n], dy = rays_d[3 * (size_t)n + 1], dz = rays_d[3 * (size_t)n + 2];
    const float A = dx * dx + dy * dy + dz * dz;
    const float Bh = ox * dx + oy * dy + oz * dz;
    const float Cc = ox * ox + oy * oy + oz * oz - radius * radius;
    const float t = (-Bh + sqrtf(Bh * Bh - A * Cc)) / A;
    const float x = ox + t * dx, y = oy + t * dy, z = oz + t * dz;
    const float theta = atan2f(sqrtf(x * x + z * z), y);
    const float phi = atan2f(z, x);
    coords[2 * (size_t)n] = 2 * theta * kRPi - 1;
    coords[2 * (size_t)n + 1] = phi * kRPi;
}

__global__ void __launch_bounds__(256) k_morton3D(const int *__restrict__ coords, uint32_t N, int *__restrict__ indices) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    indices[n] = (int)morton3D_1((uint32_t)coords[3 * (size_t)n], (uint32_t)coords[3 * (size_t)n + 1], (uint32_t)coords[3 * (size_t)n + 2]);
}

__global__ void __launch_bounds__(256) k_morton3D_invert(const int *__restrict__ indices, uint32_t N, int *__restrict__ coords) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    const int ind = indices[n];
    coords[3 * (size_t)n] = (int)compact_bits((uint32_t)(ind >> 0));
    coords[3 * (size_t)n + 1] = (int)compact_bits((uint32_t)(ind >> 1));
    coords[3 * (size_t)n + 2] = (int)compact_bits((uint32_t)(ind >> 2));
}

// mark_untrained_grid (nerf/renderer.py:433-496) as one kernel instead of a 5-deep Python loop over cell blocks, cascades
// and camera batches: one thread per cell (c, Morton index n) walks the camera list (staged 64 poses at a time in shared
// memory) and counts the cameras whose frustum, widened by one cell, contains the cell centre; cells no camera sees are
// set to -1 in the density grid (they never become occupied: update_extra_state only touches cells >= 0).
// Arithmetic = the torch expressions of the reference, one rounding per op (no contraction); the 3x3 product is summed
// ((c0 R0j + c1 R1j) + c2 R2j).  poses: [B,4,4] row-major camera-to-world.
__global__ void __launch_bounds__(256) k_mark_untrained(const float *__restrict__ poses, uint32_t B, float kx, float ky, float bound, uint32_t C,
                                                        uint32_t H, float *__restrict__ grid, int *__restrict__ count_out) {
    __shared__ float s_pose[64][12];
    const uint32_t H3 = H * H * H;
    const uint32_t n = blockIdx.x * blockDim.x + threadIdx.x, cas = blockIdx.y;
    const float bound_c = fminf(scalbnf(1.0f, (int)cas), bound);
    const double hgs_d = (double)bound_c / (double)H;
    const float scale = (float)((double)bound_c - hgs_d), margin = (float)(hgs_d * 2.0);
    float w[3];
    {
        const uint32_t c[3] = {compact_bits(n), compact_bits(n >> 1), compact_bits(n >> 2)};
#pragma unroll
        for (int d = 0; d < 3; ++d) w[d] = __fmul_rn(__fsub_rn(__fdiv_rn(__fmul_rn(2.0f, (float)c[d]), (float)(H - 1)), 1.0f), scale);
    }
    int count = 0;
    for (uint32_t b0 = 0; b0 < B; b0 += 64) {
        const uint32_t nb = min(64u, B - b0);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < nb * 12; i += blockDim.x) s_pose[i / 12][i % 12] = poses[(size_t)(b0 + i / 12) * 16 + i % 12];
        __syncthreads();
        for (uint32_t b = 0; b < nb; ++b) {
            const float *P = s_pose[b];  // rows 0..2 of the 4x4: P[4 i + j] = R[i][j], P[4 i + 3] = t[i]
            const float d0 = __fsub_rn(w[0], P[3]), d1 = __fsub_rn(w[1], P[7]), d2 = __fsub_rn(w[2], P[11]);
            float cam[3];
#pragma unroll
            for (int j = 0; j < 3; ++j)
                cam[j] = __fadd_rn(__fadd_rn(__fmul_rn(d0, P[j]), __fmul_rn(d1, P[4 + j])), __fmul_rn(d2, P[8 + j]));
            const bool in = cam[2] > 0.0f && fabsf(cam[0]) < __fadd_rn(__fmul_rn(kx, cam[2]), margin) &&
                            fabsf(cam[1]) < __fadd_rn(__fmul_rn(ky, cam[2]), margin);
            count += in ? 1 : 0;
        }
    }
    if (n >= H3) return;
    if (count_out) count_out[(size_t)cas * H3 + n] = count;
    if (count == 0) grid[(size_t)cas * H3 + n] = -1.0f;
}

// one thread packs 32 cells (128 B in, 4 B out): 8x float4 loads, one 32-bit store.
// N4 = number of whole 4-byte groups; the (< 4) trailing bytes go through k_packbits_tail.
__global__ void __launch_bounds__(256) k_packbits32(const float4 *__restrict__ grid4, uint32_t N4, float thresh,
                                                    uint32_t *__restrict__ bitfield32) {
    const uint32_t n = threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N4) return;
    uint32_t bits = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const float4 v = __ldg(grid4 + 8 * (size_t)n + i);
        bits |= (v.x > thresh ? 1u : 0u) << (4 * i) | (v.y > thresh ? 1u : 0u) << (4 * i + 1) |
                (v.z > thresh ? 1u : 0u) << (4 * i + 2) | (v.w > thresh ? 1u : 0u) << (4 * i + 3);
    }
    bitfield32[n] = bits;  // little-endian: byte k of the word == cells 8k..8k+7, LSB first
}
__global__ void k_packbits_tail(const float *__restrict__ grid, uint32_t n_begin, uint32_t N, float thresh, uint8_t *__restrict__ bitfield) {
    const uint32_t n = n_begin + threadIdx.x + blockIdx.x * blockDim.x;
    if (n >= N) return;
    uint8_t bits = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) bits |= (grid[8 * (size_t)n + i] > thresh) ? (uint8_t)(1u << i) : 0;
    bitfield[n] = bits;
}

}  // namespace
}  // namespace ngp

#include "raymarch_train.cuh"
#include "raymarch_warp.cuh"
#include "raymarch_infer.cuh"
#include "raymarch_host.cuh"

// ==========================================================================================================
// ngp_compact_alive_static -- the same compaction with NO caller workspace and no allocation of any kind
// ==========================================================================================================
// No counterpart in the reference: added for callers that cannot allocate
// (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The chained-scan scratch
// (march_ws_t: done flag, grand total, one block total per 1024 ids) lives in a file-scope device array of a fixed,
// modest size; every GPU that loads the library holds these ~16 KB until the process exits.
//
// In HOST code `g_compact_scratch` is only a host-side stand-in for the symbol, not a device address: passing it to a
// kernel faults (or, where the GPU can reach host memory, silently uses the wrong memory).  The device address is
// looked up once per device with cudaGetSymbolAddress, with that device current, and that pointer is what the kernels get.
//
// Every call on a device shares the one array.  The two words that must start at zero are cleared on the stream at the
// start of each call; the block totals are fully rewritten by k_compact_count before k_compact_write reads them, so
// the contents a previous call left behind are never read.  Calls that OVERLAP on the GPU (two streams or two threads
// of one process on the same device) race on the array: serialise them on one stream, or use ngp_compact_alive with
// one workspace per stream.
#include <mutex>

static __device__ __align__(256) uint32_t g_compact_scratch[2 + NGP_COMPACT_STATIC_MAX_N / (128 * 8)];
static_assert(ngp::kMarchBlock * ngp::kCompactItems == 128 * 8, "NGP_COMPACT_STATIC_MAX_N assumes 1024 ids per block");

namespace ngp {
namespace {
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
march_ws_t *g_scratch_ptr[kMaxDevices] = {nullptr};

// device address of g_compact_scratch on the CURRENT device (looked up once per device)
int compact_scratch(march_ws_t **out) {
    int dev = -1;
    NGP_CHECK_CUDA(cudaGetDevice(&dev));
    NGP_CHECK_ARG(dev >= 0 && dev < kMaxDevices, "compact_alive_static: device index %d out of range", dev);
    std::lock_guard<std::mutex> l(g_scratch_mu);
    if (g_scratch_ptr[dev] == nullptr) {
        void *p = nullptr;
        NGP_CHECK_CUDA(cudaGetSymbolAddress(&p, g_compact_scratch));
        NGP_CHECK_ARG(p != nullptr, "compact_alive_static: cudaGetSymbolAddress returned a null device address");
        g_scratch_ptr[dev] = reinterpret_cast<march_ws_t *>(p);
    }
    *out = g_scratch_ptr[dev];
    return NGP_OK;
}
}  // namespace
}  // namespace ngp

extern "C" int ngp_compact_alive_static(const int *in, uint32_t n, const int *n_dev, int *out, int *n_out, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(in && out && n_out && in != out, "compact_alive_static: null or aliasing in/out");
    NGP_CHECK_ARG(n <= (uint32_t)NGP_COMPACT_STATIC_MAX_N,
                  "compact_alive_static: n = %u exceeds NGP_COMPACT_STATIC_MAX_N = %u (the fixed device scratch); use ngp_compact_alive "
                  "with a caller workspace for larger lists", n, (uint32_t)NGP_COMPACT_STATIC_MAX_N);
    march_ws_t *ws = nullptr;
    const int rc = compact_scratch(&ws);
    if (rc != NGP_OK) return rc;
    const uint32_t nb = div_round_up(n ? n : 1u, (uint32_t)(kMarchBlock * kCompactItems));
    cudaStream_t st = (cudaStream_t)stream;
    NGP_CHECK_CUDA(cudaMemsetAsync(ws, 0, 2 * sizeof(uint32_t), st));  // done flag + grand total start at zero; no allocation
    { NGP_PROF("k_compact_count", st); k_compact_count<<<nb, kMarchBlock, 0, st>>>(in, n, n_dev, n_out, ws); }
    NGP_CHECK_LAUNCH();
    { NGP_PROF("k_compact_write", st); k_compact_write<<<nb, kMarchBlock, 0, st>>>(in, n, n_dev, out, ws); }
    NGP_CHECK_LAUNCH();
    { NGP_PROF("k_compact_publish", st); k_compact_publish<<<1, 1, 0, st>>>(ws, n_out); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}
