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

// position of one point inside one level (reference :146-156); returns false when out of [0,1]^3
struct loc3_t {
    uint32_t p0, p1, p2;   // base cell
    float f0, f1, f2;      // (smooth-stepped) fractions
};
__device__ __forceinline__ bool locate3(const float *__restrict__ inputs, uint32_t b, float scale, bool align_corners,
                                        uint32_t interp, loc3_t &o) {
    const float x0 = inputs[3 * (size_t)b], x1 = inputs[3 * (size_t)b + 1], x2 = inputs[3 * (size_t)b + 2];
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
                                                       uint32_t interp, int layout, uint32_t level_begin) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const uint32_t level = level_begin + blockIdx.y;
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    T *out = outputs + (layout == NGP_LAYOUT_LBC ? ((size_t)level * B + b) * 2 : ((size_t)b * L + level) * 2);
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
                                                             bool align_corners, uint32_t interp, int layout) {
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
        T *out = outputs + (layout == NGP_LAYOUT_LBC ? ((size_t)level * B + b) * 2 : ((size_t)b * L + level) * 2);
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
                                                       uint32_t interp, int layout, uint32_t level_begin, int warp_agg) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;  // blockDim.x is a multiple of 32: warps exit whole or not at all
    const uint32_t level = level_begin + blockIdx.y;
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    T *tab = grad_grid + (size_t)lv.offset * 2;

    loc3_t o;
    const bool valid = (b < B) && locate3(inputs, b, lv.scale, align_corners, interp, o);
    if (!warp_agg) {
        if (!valid) return;
        const float2 gv = entry_io<T>::loadg(grad + (layout == NGP_LAYOUT_LBC ? ((size_t)level * B + b) * 2 : ((size_t)b * L + level) * 2));
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
        gv = entry_io<T>::loadg(grad + (layout == NGP_LAYOUT_LBC ? ((size_t)level * B + b) * 2 : ((size_t)b * L + level) * 2));
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
                                                             bool align_corners, uint32_t interp, int layout) {
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
        if (!locate3(inputs, b, lv.scale, align_corners, interp, o)) continue;
        const float2 gv = entry_io<T>::loadg(grad + (layout == NGP_LAYOUT_LBC ? ((size_t)level * B + b) * 2 : ((size_t)b * L + level) * 2));
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

}  // namespace
}  // namespace ngp
