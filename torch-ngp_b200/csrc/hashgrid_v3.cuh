// hashgrid_v3.cuh -- D=3, C=2 kernels specialised per level *kind* (included by hashgrid.cu only).
//
// ncu on the lg4 kernels (profiles/r1_nerf_train_ncu_summary.json) showed the fp16 / ray-ordered training
// case issue-bound, not memory-bound: ~380 executed instructions per (point, level) against an LTS
// throughput of 38 %.  Most of them re-derive, per corner, facts that are constant for the level:
// whether it is hashed, whether its size is a power of two, whether an index has to be wrapped.  Here the
// first threads of a block classify the block's 4 levels once (reference get_grid_index :66-84):
//
//   kind 0  dense   : all three strides fit (stride^3 <= size): index = x + y*s + z*s^2 < size, never wrapped
//   kind 1  hashed  : gridtype hash, stride^3 > size, size a power of two: index = (x ^ y*P1 ^ z*P2) & (size-1).
//                     P0 == 1, so the two x-neighbours of every corner row differ by the same bit pattern
//                     x ^ (x+1): "both in one aligned 4-entry group" is ONE test per level (x & 3 != 3).
//   kind 2  other   : tiled grids, non-power-of-two hash maps, partially strided levels -> the general
//                     index code of hashgrid_fast.cuh (same results, more instructions).
//
// The branch on the kind is block-uniform.  Everything else (thread = point x 4 levels, 32 B sector
// gathers, full-sector output rows, x-pair reductions, segmented warp reduction in the backward)
// is the lg4 design.  Arithmetic order of weights and sums is unchanged: results are bit-identical to lg4.
#pragma once

namespace ngp {
namespace {

struct lv3_t {
    float scale;
    uint32_t offset;  // entries
    uint32_t s1, s2;  // dense strides
    uint32_t mask;    // size - 1 (hashed)
    uint32_t kind;
};

__device__ __forceinline__ lv3_t make_lv3(const int *__restrict__ offsets, uint32_t level, float S, uint32_t H, uint32_t gridtype,
                                          bool align_corners) {
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
    lv3_t o;
    o.scale = lv.scale;
    o.offset = lv.offset;
    o.s1 = g.st1;
    o.s2 = g.st2;
    o.mask = g.mask;
    const uint32_t s = align_corners ? lv.resolution : lv.resolution + 1;
    if (g.hashed) o.kind = g.pow2 && lv.size >= 4 ? 1u : 2u;
    else o.kind = (g.use1 && g.use2 && (unsigned long long)g.st2 * s <= lv.size) ? 0u : 2u;
    return o;
}

// cell + fractions + the 8 trilinear weights in the reference's product order ((wx*wy)*wz)
struct cell3_t {
    uint32_t p0, p1, p2;
    float w0[4], w1[4];  // row k = by + 2*bz: weight of the x and the x+1 corner
};
__device__ __forceinline__ void locate_cell3(float x0, float x1, float x2, float scale, float off, uint32_t interp, cell3_t &c) {
    float q0 = x0 * scale + off, q1 = x1 * scale + off, q2 = x2 * scale + off;
    const float fl0 = floorf(q0), fl1 = floorf(q1), fl2 = floorf(q2);
    c.p0 = (uint32_t)fl0; c.p1 = (uint32_t)fl1; c.p2 = (uint32_t)fl2;
    q0 -= fl0; q1 -= fl1; q2 -= fl2;  // == (float)p for p >= 0
    if (interp == 1) { q0 = smoothstep_f(q0); q1 = smoothstep_f(q1); q2 = smoothstep_f(q2); }
    const float a0 = 1 - q0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float wy = (k & 1) ? q1 : 1 - q1, wz = (k >> 1) ? q2 : 1 - q2;
        c.w0[k] = (a0 * wy) * wz;
        c.w1[k] = (q0 * wy) * wz;
    }
}

// ---- group access: 4 consecutive entries (one 32 B sector of an f32 table, 16 B of an f16 table) ----
struct grp_f32 { float v[8]; };
struct grp_f16 { uint4 u; };
__device__ __forceinline__ void load_grp(const float *tab, uint32_t e4, grp_f32 &g) { group_io<float>::load4(tab, e4, g.v); }
__device__ __forceinline__ void load_grp(const __half *tab, uint32_t e4, grp_f16 &g) {
    g.u = __ldg(reinterpret_cast<const uint4 *>(tab + 2 * (size_t)e4));
}
__device__ __forceinline__ float2 pick_grp(const grp_f32 &g, uint32_t j) { return pick4(g.v, j); }
__device__ __forceinline__ float2 pick_grp(const grp_f16 &g, uint32_t j) { return pick_word_h2(g.u, j); }
template <typename T> struct grp_of;
template <> struct grp_of<float> { using type = grp_f32; };
template <> struct grp_of<__half> { using type = grp_f16; };

__device__ __forceinline__ float2 blend8(const cell3_t &c, const float2 v0[4], const float2 v1[4]) {
    float a = 0, b = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // accumulation order of gather8 / the generic kernel
        a += c.w0[k] * v0[k].x; b += c.w0[k] * v0[k].y;
        a += c.w1[k] * v1[k].x; b += c.w1[k] * v1[k].y;
    }
    return make_float2(a, b);
}

constexpr uint32_t kP1 = 2654435761u, kP2 = 805459861u;

// the 8 entry indices of a cell, by kind (block-uniform branch in the callers)
__device__ __forceinline__ void indices_hashed(const lv3_t &lv, const cell3_t &c, uint32_t i0[4], uint32_t i1[4]) {
    const uint32_t hy0 = c.p1 * kP1, hy1 = hy0 + kP1, hz0 = c.p2 * kP2, hz1 = hz0 + kP2;
    const uint32_t yz[4] = {hy0 ^ hz0, hy1 ^ hz0, hy0 ^ hz1, hy1 ^ hz1};
    const uint32_t xb = c.p0 + 1;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        i0[k] = (c.p0 ^ yz[k]) & lv.mask;
        i1[k] = (xb ^ yz[k]) & lv.mask;
    }
}
__device__ __forceinline__ void indices_dense(const lv3_t &lv, const cell3_t &c, uint32_t i0[4], uint32_t i1[4]) {
    const uint32_t base = c.p0 + c.p1 * lv.s1 + c.p2 * lv.s2;
    i0[0] = base; i0[1] = base + lv.s1; i0[2] = base + lv.s2; i0[3] = base + lv.s1 + lv.s2;
#pragma unroll
    for (int k = 0; k < 4; ++k) i1[k] = i0[k] + 1;
}
__device__ __forceinline__ void indices_other(const int *__restrict__ offsets, uint32_t level, float S, uint32_t H, uint32_t gridtype,
                                           bool align_corners, const cell3_t &c, uint32_t i0[4], uint32_t i1[4]) {
    const level_t lv = level_consts(offsets, level, S, H);
    const geo3_t g = make_geo3(lv, gridtype, align_corners);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t yz = yz_term3(g, c.p1 + (k & 1), c.p2 + (k >> 1));
        i0[k] = index3(g, c.p0, yz);
        i1[k] = index3(g, c.p0 + 1, yz);
    }
}

template <typename T>
__device__ __forceinline__ float2 gather_rows(const T *__restrict__ tab, const cell3_t &c, const uint32_t i0[4], const uint32_t i1[4]) {
    float2 v0[4], v1[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        if ((i0[k] >> 2) == (i1[k] >> 2)) {
            typename grp_of<T>::type g;
            load_grp(tab, i0[k] & ~3u, g);
            v0[k] = pick_grp(g, i0[k] & 3u);
            v1[k] = pick_grp(g, i1[k] & 3u);
        } else {
            v0[k] = entry_io<T>::load1_nc(tab, i0[k]);
            v1[k] = entry_io<T>::load1_nc(tab, i1[k]);
        }
    }
    return blend8(c, v0, v1);
}

// hashed levels: same-group is a property of x alone
template <typename T>
__device__ __forceinline__ float2 gather_hashed(const T *__restrict__ tab, const lv3_t &lv, const cell3_t &c) {
    uint32_t i0[4], i1[4];
    indices_hashed(lv, c, i0, i1);
    float2 v0[4], v1[4];
    if ((c.p0 & 3u) != 3u) {
        typename grp_of<T>::type g[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) load_grp(tab, i0[k] & ~3u, g[k]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v0[k] = pick_grp(g[k], i0[k] & 3u);
            v1[k] = pick_grp(g[k], i1[k] & 3u);
        }
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            v0[k] = entry_io<T>::load1_nc(tab, i0[k]);
            v1[k] = entry_io<T>::load1_nc(tab, i1[k]);
        }
    }
    return blend8(c, v0, v1);
}

// Thread -> (point, first level) mapping.
//   MIX = false: grid = (ceil(B/256), L/4); a block works on one group of 4 levels, the level groups run one after
//                the other (blockIdx.y-major scheduling).
//   MIX = true : grid = ceil(B / (32 * 8/NG)), NG = L/4 in {1,2,4,8}; warp w of a block works on level group w % NG.
//                Every SM then always holds a mix of coarse-level warps (instruction-issue bound on ray-ordered
//                samples: whole warps share a few cells) and fine-level warps (bound by the one L1-miss request per
//                clock an SM can send to L2), so both limits are worked on at the same time instead of in turn.
template <bool MIX>
__device__ __forceinline__ void map_thread(uint32_t L, uint32_t &b, uint32_t &l0, uint32_t lg_first = 0, uint32_t ng_call = 0) {
    if (MIX) {
        // ng_call: level groups of THIS launch (a call over a level range, ngp_call_opts.level_begin / level_end); 0 = all of them
        const uint32_t ng = ng_call ? ng_call : L >> 2, warp = threadIdx.x >> 5;
        const uint32_t grp = warp & (ng - 1u), sub = warp / ng;
        b = blockIdx.x * (256u / ng) + sub * 32u + (threadIdx.x & 31u);
        l0 = (grp + lg_first) * 4u;
    } else {
        b = blockIdx.x * blockDim.x + threadIdx.x;
        l0 = (blockIdx.y + lg_first) * 4u;  // lg_first: a call over a level range (ngp_call_opts.level_begin)
    }
}
template <bool MIX>
__device__ __forceinline__ const lv3_t *block_levels(lv3_t *s_lv, const int *__restrict__ offsets, uint32_t L, uint32_t l0, float S,
                                                     uint32_t H, uint32_t gridtype, bool align_corners) {
    if (MIX) {
        if (threadIdx.x < L) s_lv[threadIdx.x] = make_lv3(offsets, threadIdx.x, S, H, gridtype, align_corners);
    } else {
        if (threadIdx.x < 4) s_lv[threadIdx.x] = make_lv3(offsets, l0 + threadIdx.x, S, H, gridtype, align_corners);
    }
    __syncthreads();
    return MIX ? s_lv + l0 : s_lv;
}

// forward: L % 4 == 0, table and output 32 B aligned (host-checked)
template <typename T, int MINB, bool MIX>
__global__ void __launch_bounds__(256, MINB) k_grid_fwd_v3(const float *__restrict__ inputs, const T *__restrict__ grid,
                                                     const int *__restrict__ offsets, T *__restrict__ outputs, uint32_t B, uint32_t L,
                                                     float S, uint32_t H, uint32_t gridtype, bool align_corners, uint32_t interp,
                                                     int layout, float in_bound, const int *__restrict__ limit) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    uint32_t b, l0;
    map_thread<MIX>(L, b, l0);
    __shared__ lv3_t s_lv[MIX ? 32 : 4];
    const lv3_t *my_lv = block_levels<MIX>(s_lv, offsets, L, l0, S, H, gridtype, align_corners);
    if (b >= B) return;
    float x0 = inputs[3 * (size_t)b], x1 = inputs[3 * (size_t)b + 1], x2 = inputs[3 * (size_t)b + 2];
    if (in_bound > 0.f) { x0 = to_unit(x0, in_bound); x1 = to_unit(x1, in_bound); x2 = to_unit(x2, in_bound); }
    const bool oob = (x0 < 0 || x0 > 1 || x1 < 0 || x1 > 1 || x2 < 0 || x2 > 1);
    const float off = align_corners ? 0.0f : 0.5f;
    float2 res[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        res[k] = make_float2(0.f, 0.f);
        if (oob) continue;
        const lv3_t lv = my_lv[k];
        cell3_t c;
        locate_cell3(x0, x1, x2, lv.scale, off, interp, c);
        const T *tab = grid + (size_t)lv.offset * 2;
        if (lv.kind == 1u) {
            res[k] = gather_hashed<T>(tab, lv, c);
        } else {
            uint32_t i0[4], i1[4];
            if (lv.kind == 0u) indices_dense(lv, c, i0, i1);
            else indices_other(offsets, l0 + k, S, H, gridtype, align_corners, c, i0, i1);
            res[k] = gather_rows<T>(tab, c, i0, i1);
        }
    }
    if (layout == NGP_LAYOUT_BLC) {
        store_row4(outputs + ((size_t)b * L + l0) * 2, res);
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) entry_io<T>::store1(outputs + ((size_t)(l0 + k) * Bs + b) * 2, res[k].x, res[k].y);
    }
}

// ------------------------------------------------------------------------------------------
// backward
// ------------------------------------------------------------------------------------------
// fp16 table: both x-neighbours inside one aligned 4-entry group (16 B) -> ONE red.v4.f16x2 (+0 on the two entries
// that are not ours) instead of two 4 B reductions.  The L2 atomic units are the limit of this kernel and work per
// request, not per byte.
__device__ __forceinline__ void red_group_h(__half *tab, uint32_t i0, uint32_t i1, float a0, float b0, float a1, float b1) {
    const uint32_t u0 = pack_h2(a0, b0), u1 = pack_h2(a1, b1);
    const uint32_t j0 = i0 & 3u, j1 = i1 & 3u;
    const uint32_t w0 = j0 == 0u ? u0 : (j1 == 0u ? u1 : 0u), w1 = j0 == 1u ? u0 : (j1 == 1u ? u1 : 0u);
    const uint32_t w2 = j0 == 2u ? u0 : (j1 == 2u ? u1 : 0u), w3 = j0 == 3u ? u0 : (j1 == 3u ? u1 : 0u);
    asm volatile("red.global.add.noftz.v4.f16x2 [%0], {%1, %2, %3, %4};" ::"l"(tab + 2 * (size_t)(i0 & ~3u)), "r"(w0), "r"(w1), "r"(w2), "r"(w3)
                 : "memory");
}

template <typename T>
__device__ __forceinline__ void red_rows(T *__restrict__ tab, const uint32_t i0[4], const uint32_t i1[4], const float s[16], int red4) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if constexpr (sizeof(T) == 2) {
            if (red4 && (i0[q] >> 2) == (i1[q] >> 2) && (i0[q] ^ i1[q]) != 1u) {
                red_group_h(tab, i0[q], i1[q], s[4 * q + 0], s[4 * q + 1], s[4 * q + 2], s[4 * q + 3]);
                continue;
            }
        }
        if ((i0[q] ^ i1[q]) == 1u) {
            if (i0[q] & 1u) entry_io<T>::red2(tab, i1[q], s[4 * q + 2], s[4 * q + 3], s[4 * q + 0], s[4 * q + 1]);
            else entry_io<T>::red2(tab, i0[q], s[4 * q + 0], s[4 * q + 1], s[4 * q + 2], s[4 * q + 3]);
        } else {
            entry_io<T>::red1(tab, i0[q], s[4 * q + 0], s[4 * q + 1]);
            entry_io<T>::red1(tab, i1[q], s[4 * q + 2], s[4 * q + 3]);
        }
    }
}

// Segmented warp reduction over runs of equal cells (see k_grid_bwd_d3c2_lg4), with two changes:
// the shuffle rounds stop at the longest run of the warp (REDUX max) instead of always running five,
// and the index arithmetic is the per-kind code above.
template <typename T, int MINB, bool MIX, bool UNROLL>
__global__ void __launch_bounds__(256, MINB) k_grid_bwd_v3(const T *__restrict__ grad, const float *__restrict__ inputs,
                                                     const int *__restrict__ offsets, T *__restrict__ grad_grid, uint32_t B, uint32_t L,
                                                     float S, uint32_t H, uint32_t gridtype, bool align_corners, uint32_t interp,
                                                     int layout, int warp_agg, int red4, uint32_t skip, float in_bound, const int *__restrict__ limit,
                                                     uint32_t lg_first, uint32_t ng_call) {
    const uint32_t Bs = B;  // LBC row stride: the caller's B, not the device-side limit
    B = eff_batch(B, limit);
    uint32_t b, l0;
    map_thread<MIX>(L, b, l0, lg_first, ng_call);
    __shared__ lv3_t s_lv[MIX ? 32 : 4];
    const lv3_t *my_lv = block_levels<MIX>(s_lv, offsets, L, l0, S, H, gridtype, align_corners);
    float x0 = 0.f, x1 = 0.f, x2 = 0.f;
    if (b < B) { x0 = inputs[3 * (size_t)b]; x1 = inputs[3 * (size_t)b + 1]; x2 = inputs[3 * (size_t)b + 2]; }
    if (in_bound > 0.f) { x0 = to_unit(x0, in_bound); x1 = to_unit(x1, in_bound); x2 = to_unit(x2, in_bound); }
    const bool inside = (b < B) && !(x0 < 0 || x0 > 1 || x1 < 0 || x1 > 1 || x2 < 0 || x2 > 1);
    if (__ballot_sync(0xffffffffu, inside) == 0) return;
    const float off = align_corners ? 0.0f : 0.5f;
    const uint32_t lane = threadIdx.x & 31u;
    float2 gv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) gv[k] = make_float2(0.f, 0.f);
    if (inside) {
        if (layout == NGP_LAYOUT_BLC) {
            load_row4(grad + ((size_t)b * L + l0) * 2, gv);
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) gv[k] = entry_io<T>::loadg(grad + ((size_t)(l0 + k) * Bs + b) * 2);
        }
    }
    // UNROLL = false: the body (three index kinds, sixteen-value segmented reduction, reduction variants) is ~2 k instructions;
    // four copies made a 131 KB kernel whose warps -- each on its own level group and kind -- thrashed the instruction cache
    // (ncu: "no_instruction" was the top stall reason).  Measured back to back in one job (option hashgrid_v3_unroll_bwd):
    // ray-ordered fp16 samples 309-323 -> 270-280 ps/sample (the path that runs the segmented reduction); unordered points
    // (configs[0], the SDF step) do not care either way.  The level's gradient is picked with selects so gv[] stays in registers.
#pragma unroll(UNROLL ? 4 : 1)
    for (int k = 0; k < 4; ++k) {
        if (l0 + k < skip) continue;  // coarse levels accumulated in shared memory by k_grid_bwd_d3c2_smem
        const lv3_t lv = my_lv[k];
        T *tab = grad_grid + (size_t)lv.offset * 2;
        const float2 gk = k == 0 ? gv[0] : (k == 1 ? gv[1] : (k == 2 ? gv[2] : gv[3]));
        // A zero upstream gradient adds nothing: drop the lane for this level.  In NeRF training every sample behind
        // the point where a ray's transmittance fell below T_thresh has an exactly zero gradient.
        const bool valid = inside && (gk.x != 0.f || gk.y != 0.f);
        if (__ballot_sync(0xffffffffu, valid) == 0) continue;
        cell3_t c;
        uint32_t i0[4], i1[4];
        locate_cell3(valid ? x0 : 0.f, valid ? x1 : 0.f, valid ? x2 : 0.f, lv.scale, off, interp, c);
        if (lv.kind == 1u) indices_hashed(lv, c, i0, i1);
        else if (lv.kind == 0u) indices_dense(lv, c, i0, i1);
        else indices_other(offsets, l0 + k, S, H, gridtype, align_corners, c, i0, i1);
        float s[16];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            s[4 * q + 0] = c.w0[q] * gk.x; s[4 * q + 1] = c.w0[q] * gk.y;
            s[4 * q + 2] = c.w1[q] * gk.x; s[4 * q + 3] = c.w1[q] * gk.y;
        }
        if (!warp_agg) {
            if (valid) red_rows<T>(tab, i0, i1, s, red4);
            continue;
        }
        unsigned long long key = (unsigned long long)c.p0 | ((unsigned long long)c.p1 << 21) | ((unsigned long long)c.p2 << 42);
        if (!valid) key = ~0ull - lane;
        const unsigned long long prev = __shfl_up_sync(0xffffffffu, key, 1);
        const uint32_t heads = __ballot_sync(0xffffffffu, lane == 0 || prev != key);
        if (__popc(heads) > 24) {  // (almost) no runs: every lane reduces for itself
            // red4 == 2: group-wide reductions only when the lanes of the warp do not meet in the same groups (random
            // points).  On ray-ordered samples neighbouring lanes hit the same 16 B group and the wider reduction,
            // which touches all four entries, makes the L2 atomic unit serialise more of them (measured: -9 % time on
            // the SDF step's random points, +20 % on NeRF samples).
            int r4 = red4;
            if (red4 == 2) {
                const uint32_t grp = i0[0] >> 2, pg = __shfl_up_sync(0xffffffffu, grp, 1);
                r4 = __popc(__ballot_sync(0xffffffffu, lane != 0 && pg == grp)) < 2;
            }
            if (valid) red_rows<T>(tab, i0, i1, s, r4);
            continue;
        }
        const int start = 31 - __clz((int)(heads & (0xffffffffu >> (31u - lane))));  // first lane of my run
        const int maxrun = __reduce_max_sync(0xffffffffu, (int)lane - start + 1);
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            if (d < maxrun) {
                const bool take = (int)lane - d >= start;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float u = __shfl_up_sync(0xffffffffu, s[j], d);
                    if (take) s[j] += u;
                }
            }
        }
        const bool tail = lane == 31u || ((heads >> (lane + 1u)) & 1u);
        if (tail && valid) red_rows<T>(tab, i0, i1, s, red4 == 1);
    }
}

}  // namespace
}  // namespace ngp
