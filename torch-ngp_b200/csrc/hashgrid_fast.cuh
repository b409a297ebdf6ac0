// hashgrid_fast.cuh -- D=3, C=2 fast paths (the Instant-NGP L=16, F=2 configuration).
// Included by hashgrid.cu only.
#pragma once

namespace ngp {
namespace {

// ------------------------------------------------------------------------------------------
// level geometry for D=3: replays the stride loop of the reference's get_grid_index (:68-81)
// once per thread instead of once per corner.
// ------------------------------------------------------------------------------------------
struct geo3_t {
    uint32_t st1, st2, size, mask;
    bool use1, use2, hashed, pow2;
};

__device__ __forceinline__ geo3_t make_geo3(const level_t &lv, uint32_t gridtype, bool align_corners) {
    geo3_t g;
    const uint32_t s = align_corners ? lv.resolution : lv.resolution + 1;
    g.size = lv.size;
    uint32_t stride = s;  // d == 0 always runs (1 <= size)
    g.st1 = stride;
    g.use1 = stride <= lv.size;
    if (g.use1) stride *= s;
    g.st2 = stride;
    g.use2 = g.use1 && stride <= lv.size;
    if (g.use2) stride *= s;
    g.hashed = (gridtype == 0) && (stride > lv.size);
    g.pow2 = (lv.size & (lv.size - 1)) == 0;
    g.mask = lv.size - 1;
    return g;
}

__device__ __forceinline__ uint32_t wrap3(const geo3_t &g, uint32_t i) {
    if (g.pow2) return i & g.mask;
    return i >= g.size ? i % g.size : i;
}
// contribution of (y, z) to the index; combined with x by ^ (hashed) or + (dense/tiled)
__device__ __forceinline__ uint32_t yz_term3(const geo3_t &g, uint32_t y, uint32_t z) {
    if (g.hashed) return (y * 2654435761u) ^ (z * 805459861u);
    return (g.use1 ? y * g.st1 : 0u) + (g.use2 ? z * g.st2 : 0u);
}
__device__ __forceinline__ uint32_t index3(const geo3_t &g, uint32_t x, uint32_t yz) {
    return wrap3(g, g.hashed ? (x ^ yz) : (x + yz));
}

// ------------------------------------------------------------------------------------------
// entry (2 channels) / aligned entry-pair access for f32 and f16 tables
// ------------------------------------------------------------------------------------------
template <typename T> struct entry_io;
template <> struct entry_io<float> {
    // `base` may point to global or shared memory (generic addressing)
    static __device__ __forceinline__ float2 load1(const float *base, uint32_t e) {
        return *reinterpret_cast<const float2 *>(base + 2 * (size_t)e);
    }
    // aligned pair {e & ~1, e | 1}
    static __device__ __forceinline__ void load2(const float *base, uint32_t e_even, float2 &lo, float2 &hi) {
        const float4 v = *reinterpret_cast<const float4 *>(base + 2 * (size_t)e_even);
        lo = make_float2(v.x, v.y);
        hi = make_float2(v.z, v.w);
    }
    static __device__ __forceinline__ float2 load1_nc(const float *base, uint32_t e) {
        return __ldg(reinterpret_cast<const float2 *>(base + 2 * (size_t)e));
    }
    static __device__ __forceinline__ void load2_nc(const float *base, uint32_t e_even, float2 &lo, float2 &hi) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(base + 2 * (size_t)e_even));
        lo = make_float2(v.x, v.y);
        hi = make_float2(v.z, v.w);
    }
    static __device__ __forceinline__ void store1(float *p, float a, float b) { *reinterpret_cast<float2 *>(p) = make_float2(a, b); }
    static __device__ __forceinline__ float2 loadg(const float *p) { return __ldg(reinterpret_cast<const float2 *>(p)); }
    static __device__ __forceinline__ void red1(float *base, uint32_t e, float a, float b) { red_add_v2(base + 2 * (size_t)e, a, b); }
    static __device__ __forceinline__ void red2(float *base, uint32_t e_even, float a, float b, float c, float d) {
        red_add_v4(base + 2 * (size_t)e_even, a, b, c, d);
    }
};
template <> struct entry_io<__half> {
    static __device__ __forceinline__ float2 load1(const __half *base, uint32_t e) {
        return __half22float2(*reinterpret_cast<const __half2 *>(base + 2 * (size_t)e));
    }
    static __device__ __forceinline__ void load2(const __half *base, uint32_t e_even, float2 &lo, float2 &hi) {
        const uint2 v = *reinterpret_cast<const uint2 *>(base + 2 * (size_t)e_even);
        lo = __half22float2(*reinterpret_cast<const __half2 *>(&v.x));
        hi = __half22float2(*reinterpret_cast<const __half2 *>(&v.y));
    }
    static __device__ __forceinline__ float2 load1_nc(const __half *base, uint32_t e) {
        const uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(base + 2 * (size_t)e));
        return __half22float2(*reinterpret_cast<const __half2 *>(&v));
    }
    static __device__ __forceinline__ void load2_nc(const __half *base, uint32_t e_even, float2 &lo, float2 &hi) {
        const uint2 v = __ldg(reinterpret_cast<const uint2 *>(base + 2 * (size_t)e_even));
        lo = __half22float2(*reinterpret_cast<const __half2 *>(&v.x));
        hi = __half22float2(*reinterpret_cast<const __half2 *>(&v.y));
    }
    static __device__ __forceinline__ void store1(__half *p, float a, float b) { *reinterpret_cast<__half2 *>(p) = __floats2half2_rn(a, b); }
    static __device__ __forceinline__ float2 loadg(const __half *p) {
        const uint32_t v = __ldg(reinterpret_cast<const uint32_t *>(p));
        return __half22float2(*reinterpret_cast<const __half2 *>(&v));
    }
    static __device__ __forceinline__ void red1(__half *base, uint32_t e, float a, float b) { red_add_h2(base + 2 * (size_t)e, a, b); }
    static __device__ __forceinline__ void red2(__half *base, uint32_t e_even, float a, float b, float c, float d) {
        red_add_h2x2(base + 2 * (size_t)e_even, a, b, c, d);
    }
};

// (x + bound) / (2 bound): gridencoder/grid.py:807, the arithmetic of k_nerf_x01 (IEEE division, no contraction)
__device__ __forceinline__ float to_unit(float x, float bound) { return __fdiv_rn(__fadd_rn(x, bound), __fmul_rn(2.0f, bound)); }

// position of one point inside one level (reference :146-156); returns false when out of [0,1]^3
struct loc3_t {
    uint32_t p0, p1, p2;   // base cell
    float f0, f1, f2;      // (smooth-stepped) fractions
};
__device__ __forceinline__ bool locate3(const float *__restrict__ inputs, uint32_t b, float scale, bool align_corners,
                                        uint32_t interp, loc3_t &o, float in_bound = 0.f) {
    float x0 = inputs[3 * (size_t)b], x1 = inputs[3 * (size_t)b + 1], x2 = inputs[3 * (size_t)b + 2];
    if (in_bound > 0.f) { x0 = to_unit(x0, in_bound); x1 = to_unit(x1, in_bound); x2 = to_unit(x2, in_bound); }
    if (x0 < 0 || x0 > 1 || x1 < 0 || x1 > 1 || x2 < 0 || x2 > 1) return false;
    const float off = align_corners ? 0.0f : 0.5f;
    float q0 = x0 * scale + off, q1 = x1 * scale + off, q2 = x2 * scale + off;
    o.p0 = (uint32_t)floorf(q0); o.p1 = (uint32_t)floorf(q1); o.p2 = (uint32_t)floorf(q2);
    q0 -= (float)o.p0; q1 -= (float)o.p1; q2 -= (float)o.p2;
    if (interp == 1) { q0 = smoothstep_f(q0); q1 = smoothstep_f(q1); q2 = smoothstep_f(q2); }
    o.f0 = q0; o.f1 = q1; o.f2 = q2;
    return true;
}

// the 4 (y,z) corner rows: indices of the two x-neighbours and the row weight, in the reference's
// corner order idx = x + 2y + 4z, weight product order ((wx * wy) * wz).
struct rows3_t {
    uint32_t i0[4], i1[4];
    float w0[4], w1[4];
};
__device__ __forceinline__ void make_rows3(const geo3_t &g, const loc3_t &o, rows3_t &r) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t by = k & 1, bz = k >> 1;
        const uint32_t yz = yz_term3(g, o.p1 + by, o.p2 + bz);
        r.i0[k] = index3(g, o.p0, yz);
        r.i1[k] = index3(g, o.p0 + 1, yz);
        const float wy = by ? o.f1 : 1 - o.f1, wz = bz ? o.f2 : 1 - o.f2;
        r.w0[k] = ((1 - o.f0) * wy) * wz;
        r.w1[k] = (o.f0 * wy) * wz;
    }
}

// gather + blend the 8 corners.  SMEM: table slice lives in shared memory (no __ldg).
template <typename T, bool SMEM>
__device__ __forceinline__ float2 gather8(const T *__restrict__ tab, const rows3_t &r) {
    float2 v0[4], v1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if ((r.i0[k] ^ r.i1[k]) == 1u) {  // x-neighbours share an aligned pair slot: one wide access
            float2 lo, hi;
            if (SMEM) entry_io<T>::load2(tab, r.i0[k] & ~1u, lo, hi);
            else entry_io<T>::load2_nc(tab, r.i0[k] & ~1u, lo, hi);
            const bool odd = r.i0[k] & 1u;
            v0[k] = odd ? hi : lo;
            v1[k] = odd ? lo : hi;
        } else {
            if (SMEM) { v0[k] = entry_io<T>::load1(tab, r.i0[k]); v1[k] = entry_io<T>::load1(tab, r.i1[k]); }
            else { v0[k] = entry_io<T>::load1_nc(tab, r.i0[k]); v1[k] = entry_io<T>::load1_nc(tab, r.i1[k]); }
        }
    }
    float a = 0, b = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        a += r.w0[k] * v0[k].x; b += r.w0[k] * v0[k].y;
        a += r.w1[k] * v1[k].x; b += r.w1[k] * v1[k].y;
    }
    return make_float2(a, b);
}

// ------------------------------------------------------------------------------------------
// forward, levels read through L1/L2: one thread = one (point, level)
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) k_grid_fwd_d3c2(const float *__restrict__ inputs, const T *__restrict__ grid,
                                                       const int *__restrict__ offsets, T *__restrict__ outputs, uint32_t B,
                                                       uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                                                       uint32_t interp, int layout, uint32_t level_begin, const int *__restrict__ limit) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint32_t level = level_begin + blockIdx.y;
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    T *out = outputs + (layout == NGP_LAYOUT_LBC ? ((size_t)level * Bs + b) * 2 : ((size_t)b * L + level) * 2);
    loc3_t o;
    if (!locate3(inputs, b, lv.scale, align_corners, interp, o)) {
        entry_io<T>::store1(out, 0.f, 0.f);
        return;
    }
    rows3_t r;
    make_rows3(g, o, r);
    const float2 res = gather8<T, false>(grid + (size_t)lv.offset * 2, r);
    entry_io<T>::store1(out, res.x, res.y);
}

// ------------------------------------------------------------------------------------------
// forward, coarse levels resident in shared memory.  grid = (blocks_per_level, n_levels); each
// block copies its level's slice (<= smem budget) once and then serves a grid-stride share of
// the points out of shared memory (bank-conflict cost only, no L1 tag / L2 sector traffic).
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) k_grid_fwd_d3c2_smem(const float *__restrict__ inputs, const T *__restrict__ grid,
                                                             const int *__restrict__ offsets, T *__restrict__ outputs,
                                                             uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype,
                                                             bool align_corners, uint32_t interp, int layout, const int *__restrict__ limit) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T *tab = reinterpret_cast<T *>(smem_raw);
    const uint32_t level = blockIdx.y;
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    {   // slice is a multiple of 8 entries and starts on a multiple of 8 entries (host-checked): 16 B chunks
        const uint4 *src = reinterpret_cast<const uint4 *>(grid + (size_t)lv.offset * 2);
        uint4 *dst = reinterpret_cast<uint4 *>(tab);
        const uint32_t n16 = (uint32_t)(((size_t)lv.size * 2 * sizeof(T)) / 16);
        for (uint32_t i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < B; b += gridDim.x * blockDim.x) {
        T *out = outputs + (layout == NGP_LAYOUT_LBC ? ((size_t)level * Bs + b) * 2 : ((size_t)b * L + level) * 2);
        loc3_t o;
        if (!locate3(inputs, b, lv.scale, align_corners, interp, o)) {
            entry_io<T>::store1(out, 0.f, 0.f);
            continue;
        }
        rows3_t r;
        make_rows3(g, o, r);
        const float2 res = gather8<T, true>(tab, r);
        entry_io<T>::store1(out, res.x, res.y);
    }
}

// ------------------------------------------------------------------------------------------
// backward through L2 reductions: one thread = one (point, level).  If every lane of the warp
// sits in the same cell (coherent samples along a ray on coarse levels) the 32x8 same-address
// reductions collapse to 8 after a shuffle tree.
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <typename T>
__device__ __forceinline__ void scatter8(T *__restrict__ tab, const rows3_t &r, float g0, float g1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float a0 = r.w0[k] * g0, b0 = r.w0[k] * g1, a1 = r.w1[k] * g0, b1 = r.w1[k] * g1;
        if ((r.i0[k] ^ r.i1[k]) == 1u) {
            if (r.i0[k] & 1u) entry_io<T>::red2(tab, r.i1[k], a1, b1, a0, b0);
            else entry_io<T>::red2(tab, r.i0[k], a0, b0, a1, b1);
        } else {
            entry_io<T>::red1(tab, r.i0[k], a0, b0);
            entry_io<T>::red1(tab, r.i1[k], a1, b1);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_grid_bwd_d3c2(const T *__restrict__ grad, const float *__restrict__ inputs,
                                                       const int *__restrict__ offsets, T *__restrict__ grad_grid, uint32_t B,
                                                       uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                                                       uint32_t interp, int layout, uint32_t level_begin, int warp_agg, const int *__restrict__ limit) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;  // blockDim.x is a multiple of 32: warps exit whole or not at all
    const uint32_t level = level_begin + blockIdx.y;
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    T *tab = grad_grid + (size_t)lv.offset * 2;

    loc3_t o;
    const bool valid = (b < B) && locate3(inputs, b, lv.scale, align_corners, interp, o);
    if (!warp_agg) {
        if (!valid) return;
        const float2 gv = entry_io<T>::loadg(grad + (layout == NGP_LAYOUT_LBC ? ((size_t)level * Bs + b) * 2 : ((size_t)b * L + level) * 2));
        rows3_t r;
        make_rows3(g, o, r);
        scatter8<T>(tab, r, gv.x, gv.y);
        return;
    }
    if (__ballot_sync(0xffffffffu, valid) == 0) return;
    float2 gv = make_float2(0.f, 0.f);
    rows3_t r;
    unsigned long long key = ~0ull - threadIdx.x;  // distinct per lane when invalid
    if (valid) {
        gv = entry_io<T>::loadg(grad + (layout == NGP_LAYOUT_LBC ? ((size_t)level * Bs + b) * 2 : ((size_t)b * L + level) * 2));
        make_rows3(g, o, r);
        key = (unsigned long long)o.p0 | ((unsigned long long)o.p1 << 21) | ((unsigned long long)o.p2 << 42);
    }
    int uniform = 0;
    __match_all_sync(0xffffffffu, key, &uniform);
    if (uniform) {  // all 32 lanes valid and in one cell: identical indices, reduce the weighted gradients
        float s[16];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            s[4 * k + 0] = warp_sum(r.w0[k] * gv.x); s[4 * k + 1] = warp_sum(r.w0[k] * gv.y);
            s[4 * k + 2] = warp_sum(r.w1[k] * gv.x); s[4 * k + 3] = warp_sum(r.w1[k] * gv.y);
        }
        if ((threadIdx.x & 31) == 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if ((r.i0[k] ^ r.i1[k]) == 1u) {
                    if (r.i0[k] & 1u) entry_io<T>::red2(tab, r.i1[k], s[4 * k + 2], s[4 * k + 3], s[4 * k + 0], s[4 * k + 1]);
                    else entry_io<T>::red2(tab, r.i0[k], s[4 * k + 0], s[4 * k + 1], s[4 * k + 2], s[4 * k + 3]);
                } else {
                    entry_io<T>::red1(tab, r.i0[k], s[4 * k + 0], s[4 * k + 1]);
                    entry_io<T>::red1(tab, r.i1[k], s[4 * k + 2], s[4 * k + 3]);
                }
            }
        }
        return;
    }
    if (valid) scatter8<T>(tab, r, gv.x, gv.y);
}

// ------------------------------------------------------------------------------------------
// backward, coarse levels accumulated in shared memory (fp32), flushed once per block.
// ------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(1024) k_grid_bwd_d3c2_smem(const T *__restrict__ grad, const float *__restrict__ inputs,
                                                             const int *__restrict__ offsets, T *__restrict__ grad_grid,
                                                             uint32_t B, uint32_t L, float S, uint32_t H, uint32_t gridtype,
                                                             bool align_corners, uint32_t interp, int layout, const int *__restrict__ limit,
                                                             float in_bound = 0.f) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float *acc = reinterpret_cast<float *>(smem_raw);
    const uint32_t level = blockIdx.y;
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    const uint32_t n4 = lv.size * 2 / 4;  // size is a multiple of 8 entries (host-checked)
    for (uint32_t i = threadIdx.x; i < n4; i += blockDim.x) reinterpret_cast<float4 *>(acc)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncthreads();
    for (uint32_t b = blockIdx.x * blockDim.x + threadIdx.x; b < B; b += gridDim.x * blockDim.x) {
        loc3_t o;
        if (!locate3(inputs, b, lv.scale, align_corners, interp, o, in_bound)) continue;
        const float2 gv = entry_io<T>::loadg(grad + (layout == NGP_LAYOUT_LBC ? ((size_t)level * Bs + b) * 2 : ((size_t)b * L + level) * 2));
        rows3_t r;
        make_rows3(g, o, r);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            atomicAdd(acc + 2 * r.i0[k], r.w0[k] * gv.x); atomicAdd(acc + 2 * r.i0[k] + 1, r.w0[k] * gv.y);
            atomicAdd(acc + 2 * r.i1[k], r.w1[k] * gv.x); atomicAdd(acc + 2 * r.i1[k] + 1, r.w1[k] * gv.y);
        }
    }
    __syncthreads();
    T *tab = grad_grid + (size_t)lv.offset * 2;
    for (uint32_t i = threadIdx.x; i < n4; i += blockDim.x) {
        const float4 v = reinterpret_cast<const float4 *>(acc)[i];
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f) entry_io<T>::red2(tab, 2 * i, v.x, v.y, v.z, v.w);
    }
}

// ==========================================================================================
// v2 fast path: one thread = one point x LG consecutive levels.
//
// Measured on B200 (profiles/): both directions are bound by L2 *sector transactions* (LTS
// throughput 73-84 %, DRAM ~10 %), not by bytes -- the 46.7 MiB table lives in L2 and every
// corner row of a hashed level is a random 32 B sector.  So the kernels minimise transactions:
//   * the two x-neighbours of a corner row sit in the same 32 B sector 3 times out of 4
//     (the index differs only in its low bits): one 256-bit load (LDG.E.ENL2.256) fetches the
//     aligned 4-entry group and the two entries are selected in registers;
//   * a thread owns LG = 4 levels of its point: the point-major [B, L*C] output / upstream
//     gradient row segment is one full 32 B sector per thread instead of four partial ones,
//     the input point is read once, and 16 independent gathers are in flight per thread.
// ==========================================================================================
__device__ __forceinline__ bool locate3x(float x0, float x1, float x2, float scale, bool align_corners, uint32_t interp, loc3_t &o) {
    const float off = align_corners ? 0.0f : 0.5f;
    float q0 = x0 * scale + off, q1 = x1 * scale + off, q2 = x2 * scale + off;
    o.p0 = (uint32_t)floorf(q0); o.p1 = (uint32_t)floorf(q1); o.p2 = (uint32_t)floorf(q2);
    q0 -= (float)o.p0; q1 -= (float)o.p1; q2 -= (float)o.p2;
    if (interp == 1) { q0 = smoothstep_f(q0); q1 = smoothstep_f(q1); q2 = smoothstep_f(q2); }
    o.f0 = q0; o.f1 = q1; o.f2 = q2;
    return true;
}

__device__ __forceinline__ float2 pick4(const float v[8], uint32_t e) {
    const bool lo = e & 1u, hi = e & 2u;
    const float ax = lo ? v[2] : v[0], ay = lo ? v[3] : v[1];
    const float bx = lo ? v[6] : v[4], by = lo ? v[7] : v[5];
    return make_float2(hi ? bx : ax, hi ? by : ay);
}

template <typename T> struct group_io;
template <> struct group_io<float> {
    // aligned group of 4 entries (32 B) starting at entry e4 (multiple of 4)
    static __device__ __forceinline__ void load4(const float *base, uint32_t e4, float v[8]) {
        asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                     : "l"(base + 2 * (size_t)e4));
    }
};
template <> struct group_io<__half> {
    static __device__ __forceinline__ void load4(const __half *base, uint32_t e4, float v[8]) {
        const uint4 u = __ldg(reinterpret_cast<const uint4 *>(base + 2 * (size_t)e4));
        const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
        const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)), d = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
        v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = c.x; v[5] = c.y; v[6] = d.x; v[7] = d.y;
    }
};

__device__ __forceinline__ float2 pick_word_h2(const uint4 &u, uint32_t e) {
    const uint32_t lo = (e & 1u) ? u.y : u.x, hi = (e & 1u) ? u.w : u.z;
    const uint32_t w = (e & 2u) ? hi : lo;
    return __half22float2(*reinterpret_cast<const __half2 *>(&w));
}
template <typename T>
__device__ __forceinline__ float2 gather8_v2(const T *__restrict__ tab, const rows3_t &r) {
    float2 v0[4], v1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if ((r.i0[k] >> 2) == (r.i1[k] >> 2)) {  // same 4-entry group: one sector-wide access
            if constexpr (sizeof(T) == 2) {  // fp16: select the 32-bit entry first, convert only what is used
                const uint4 u = __ldg(reinterpret_cast<const uint4 *>(tab + 2 * (size_t)(r.i0[k] & ~3u)));
                v0[k] = pick_word_h2(u, r.i0[k] & 3u);
                v1[k] = pick_word_h2(u, r.i1[k] & 3u);
            } else {
                float v[8];
                group_io<T>::load4(tab, r.i0[k] & ~3u, v);
                v0[k] = pick4(v, r.i0[k] & 3u);
                v1[k] = pick4(v, r.i1[k] & 3u);
            }
        } else {
            v0[k] = entry_io<T>::load1_nc(tab, r.i0[k]);
            v1[k] = entry_io<T>::load1_nc(tab, r.i1[k]);
        }
    }
    float a = 0, b = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // same accumulation order as gather8 / the generic kernel
        a += r.w0[k] * v0[k].x; b += r.w0[k] * v0[k].y;
        a += r.w1[k] * v1[k].x; b += r.w1[k] * v1[k].y;
    }
    return make_float2(a, b);
}

__device__ __forceinline__ void store_row4(float *p, const float2 r[4]) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r[0].x), "f"(r[0].y), "f"(r[1].x), "f"(r[1].y),
                 "f"(r[2].x), "f"(r[2].y), "f"(r[3].x), "f"(r[3].y)
                 : "memory");
}
__device__ __forceinline__ void store_row4(__half *p, const float2 r[4]) {
    uint4 u;
    u.x = pack_h2(r[0].x, r[0].y); u.y = pack_h2(r[1].x, r[1].y); u.z = pack_h2(r[2].x, r[2].y); u.w = pack_h2(r[3].x, r[3].y);
    *reinterpret_cast<uint4 *>(p) = u;
}
__device__ __forceinline__ void load_row4(const float *p, float2 r[4]) {
    float v[8];
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(p));
#pragma unroll
    for (int k = 0; k < 4; ++k) r[k] = make_float2(v[2 * k], v[2 * k + 1]);
}
__device__ __forceinline__ void load_row4(const __half *p, float2 r[4]) {
    const uint4 u = __ldg(reinterpret_cast<const uint4 *>(p));
    r[0] = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)); r[1] = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
    r[2] = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)); r[3] = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
}

// forward: grid = (ceil(B/256), L/4); L % 4 == 0, table base 32 B aligned (host-checked)
template <typename T>
__global__ void __launch_bounds__(256) k_grid_fwd_d3c2_lg4(const float *__restrict__ inputs, const T *__restrict__ grid,
                                                           const int *__restrict__ offsets, T *__restrict__ outputs, uint32_t B,
                                                           uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                                                           uint32_t interp, int layout, const int *__restrict__ limit) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l0 = blockIdx.y * 4;
    // the block's 4 levels: constants (exp2f, strides, hashed / dense) evaluated once, not once per thread
    __shared__ level_t s_lv[4];
    __shared__ geo3_t s_g[4];
    if (threadIdx.x < 4) {
        const level_t lv = level_consts(offsets, l0 + threadIdx.x, S, H);
        s_lv[threadIdx.x] = lv;
        s_g[threadIdx.x] = make_geo3(lv, gridtype, align_corners);
    }
    __syncthreads();
    if (b >= B) return;
    const float x0 = inputs[3 * (size_t)b], x1 = inputs[3 * (size_t)b + 1], x2 = inputs[3 * (size_t)b + 2];
    const bool oob = (x0 < 0 || x0 > 1 || x1 < 0 || x1 > 1 || x2 < 0 || x2 > 1);
    float2 res[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        res[k] = make_float2(0.f, 0.f);
        if (oob) continue;
        const level_t lv = s_lv[k];
        const geo3_t g = s_g[k];
        loc3_t o;
        locate3x(x0, x1, x2, lv.scale, align_corners, interp, o);
        rows3_t r;
        make_rows3(g, o, r);
        res[k] = gather8_v2<T>(grid + (size_t)lv.offset * 2, r);
    }
    if (layout == NGP_LAYOUT_BLC) {
        store_row4(outputs + ((size_t)b * L + l0) * 2, res);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) entry_io<T>::store1(outputs + ((size_t)(l0 + k) * Bs + b) * 2, res[k].x, res[k].y);
    }
}

// backward: same mapping; x-neighbours sharing an aligned pair go out as one 16 B (f32) / 8 B (f16) reduction
template <typename T>
__global__ void __launch_bounds__(256) k_grid_bwd_d3c2_lg4(const T *__restrict__ grad, const float *__restrict__ inputs,
                                                           const int *__restrict__ offsets, T *__restrict__ grad_grid, uint32_t B,
                                                           uint32_t L, float S, uint32_t H, uint32_t gridtype, bool align_corners,
                                                           uint32_t interp, int layout, int warp_agg, const int *__restrict__ limit) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l0 = blockIdx.y * 4;
    __shared__ level_t s_lv[4];
    __shared__ geo3_t s_g[4];
    if (threadIdx.x < 4) {
        const level_t lv = level_consts(offsets, l0 + threadIdx.x, S, H);
        s_lv[threadIdx.x] = lv;
        s_g[threadIdx.x] = make_geo3(lv, gridtype, align_corners);
    }
    __syncthreads();
    float x0 = 0.f, x1 = 0.f, x2 = 0.f;
    if (b < B) { x0 = inputs[3 * (size_t)b]; x1 = inputs[3 * (size_t)b + 1]; x2 = inputs[3 * (size_t)b + 2]; }
    const bool valid = (b < B) && !(x0 < 0 || x0 > 1 || x1 < 0 || x1 > 1 || x2 < 0 || x2 > 1);
    if (__ballot_sync(0xffffffffu, valid) == 0) return;
    float2 gv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) gv[k] = make_float2(0.f, 0.f);
    if (valid) {
        if (layout == NGP_LAYOUT_BLC) {
            load_row4(grad + ((size_t)b * L + l0) * 2, gv);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) gv[k] = entry_io<T>::loadg(grad + ((size_t)(l0 + k) * Bs + b) * 2);
        }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const level_t lv = s_lv[k];
        const geo3_t g = s_g[k];
        T *tab = grad_grid + (size_t)lv.offset * 2;
        loc3_t o;
        rows3_t r;
        unsigned long long key = ~0ull - threadIdx.x;
        if (valid) {
            locate3x(x0, x1, x2, lv.scale, align_corners, interp, o);
            make_rows3(g, o, r);
            key = (unsigned long long)o.p0 | ((unsigned long long)o.p1 << 21) | ((unsigned long long)o.p2 << 42);
        }
        // Segmented warp reduction.  Lanes are consecutive points; samples marched along a ray stay in one cell of a
        // coarse level for many steps (cell / dt ~ 37 samples at resolution 16, ~2 at 256), so equal keys form runs of
        // adjacent lanes.  Each run is summed with a segmented shuffle scan and its last lane issues ONE set of
        // reductions: the L2 atomic units see 1/run-length of the traffic (they, not bytes, bound this kernel), and
        // the fp16 table gradient gets one rounding per run instead of one per sample.
        uint32_t heads = 0xffffffffu;
        const uint32_t lane = threadIdx.x & 31u;
        if (warp_agg) {
            const unsigned long long prev = __shfl_up_sync(0xffffffffu, key, 1);
            heads = __ballot_sync(0xffffffffu, lane == 0 || prev != key);
        }
        if (__popc(heads) <= 20) {
            const int start = 31 - __clz((int)(heads & (0xffffffffu >> (31u - lane))));  // first lane of my run
            float s[16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                s[4 * q + 0] = valid ? r.w0[q] * gv[k].x : 0.f; s[4 * q + 1] = valid ? r.w0[q] * gv[k].y : 0.f;
                s[4 * q + 2] = valid ? r.w1[q] * gv[k].x : 0.f; s[4 * q + 3] = valid ? r.w1[q] * gv[k].y : 0.f;
            }
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const bool take = (int)lane - d >= start;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float u = __shfl_up_sync(0xffffffffu, s[j], d);
                    if (take) s[j] += u;
                }
            }
            const bool tail = lane == 31u || ((heads >> (lane + 1u)) & 1u);
            if (tail && valid) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    if ((r.i0[q] ^ r.i1[q]) == 1u) {
                        if (r.i0[q] & 1u) entry_io<T>::red2(tab, r.i1[q], s[4 * q + 2], s[4 * q + 3], s[4 * q + 0], s[4 * q + 1]);
                        else entry_io<T>::red2(tab, r.i0[q], s[4 * q + 0], s[4 * q + 1], s[4 * q + 2], s[4 * q + 3]);
                    } else {
                        entry_io<T>::red1(tab, r.i0[q], s[4 * q + 0], s[4 * q + 1]);
                        entry_io<T>::red1(tab, r.i1[q], s[4 * q + 2], s[4 * q + 3]);
                    }
                }
            }
        } else if (valid) {
            scatter8<T>(tab, r, gv[k].x, gv[k].y);
        }
    }
}

}  // namespace
}  // namespace ngp
