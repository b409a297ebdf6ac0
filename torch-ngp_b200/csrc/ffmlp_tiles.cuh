// ffmlp_tiles.cuh -- pieces shared by the FFMLP kernels (ffmlp.cu: weights resident in shared memory, hidden <= 128;
// ffmlp_wide.cu: weights streamed, hidden 256): network description, operand staging, TMEM epilogues.
#pragma once
#include "common.cuh"
#include "umma.cuh"
#include <type_traits>

namespace ngp {
namespace {

using namespace umma;

constexpr uint32_t kTileRows = 128;
constexpr uint32_t kActTileBytes = kTileRows * kTileRowBytes;  // 16 KB: 128 rows x 64 halves
constexpr uint32_t kMaxLayers = 8;                             // num_layers <= 8 (num_layers + 1 matrices)

__host__ __device__ inline uint32_t ktiles(uint32_t k) { return (k + 63u) / 64u; }
__host__ __device__ inline uint32_t tmem_cols_pow2(uint32_t n) {
    uint32_t c = 32;
    while (c < n) c <<= 1;
    return c;
}

struct mlp_t {
    uint32_t B, in, out, hid, nl;
    // weight matrix m (0..nl): rows[m] x K[m], flat offset woff[m] (elements), smem offset soff[m] (bytes)
    uint32_t rows[kMaxLayers + 1], K[kMaxLayers + 1], woff[kMaxLayers + 1], soff[kMaxLayers + 1];
    uint32_t wbytes, nweights;
    uint32_t act;  // hidden activation, the reference's codes (ffmlp.cu:22-32): 0 ReLU 1 Exponential 2 Sine 3 Sigmoid 4 Squareplus 5 Softplus 6 None
};

inline void make_mlp(mlp_t &m, uint32_t B, uint32_t in, uint32_t out, uint32_t hid, uint32_t nl) {
    m.act = 0;
    m.B = B; m.in = in; m.out = out; m.hid = hid; m.nl = nl;
    uint32_t w = 0, s = 0;
    for (uint32_t i = 0; i <= nl; ++i) {
        m.rows[i] = (i == nl) ? out : hid;
        m.K[i] = (i == 0) ? in : hid;
        m.woff[i] = w;
        m.soff[i] = s;
        w += m.rows[i] * m.K[i];
        s += ktiles(m.K[i]) * m.rows[i] * kTileRowBytes;
    }
    m.nweights = w;
    m.wbytes = (s + 1023u) & ~1023u;
}

// ---- staging ---------------------------------------------------------------------------------------
// all weight matrices -> swizzled K-major tiles ([rows x 64 halves] per 64 input columns)
template <typename WT>
__device__ __forceinline__ void stage_weights(uint8_t *sw, const WT *__restrict__ w, const mlp_t &m, uint32_t tid, uint32_t nth) {
    for (uint32_t i = 0; i <= m.nl; ++i) {
        const uint32_t rows = m.rows[i], K = m.K[i], cpr = K >> 3;
        const WT *src = w + m.woff[i];
        uint8_t *dst = sw + m.soff[i];
        for (uint32_t q = tid; q < rows * cpr; q += nth) {
            const uint32_t r = q / cpr, c8 = q - r * cpr;
            uint4 v;
            if constexpr (sizeof(WT) == 2) {
                v = __ldg(reinterpret_cast<const uint4 *>(src + (size_t)r * K) + c8);
            } else {
                const float4 a = __ldg(reinterpret_cast<const float4 *>(src + (size_t)r * K) + 2 * c8);
                const float4 b = __ldg(reinterpret_cast<const float4 *>(src + (size_t)r * K) + 2 * c8 + 1);
                const __half2 h0 = __floats2half2_rn(a.x, a.y), h1 = __floats2half2_rn(a.z, a.w);
                const __half2 h2 = __floats2half2_rn(b.x, b.y), h3 = __floats2half2_rn(b.z, b.w);
                v.x = *reinterpret_cast<const uint32_t *>(&h0); v.y = *reinterpret_cast<const uint32_t *>(&h1);
                v.z = *reinterpret_cast<const uint32_t *>(&h2); v.w = *reinterpret_cast<const uint32_t *>(&h3);
            }
            *reinterpret_cast<uint4 *>(dst + (c8 >> 3) * rows * kTileRowBytes + swz_off(r, c8 & 7u)) = v;
        }
    }
}

// rows [row0, row0+128) of a row-major fp16 [B x cols] matrix -> swizzled activation tiles, asynchronously;
// rows >= B are zero-filled (cp.async src-size 0)
// off8: first 16-byte chunk of the tile row to fill (packs two narrow operands side by side in one tile)
__device__ __forceinline__ void load_rows_async(uint8_t *tiles, const __half *__restrict__ g, uint32_t row0, uint32_t B, uint32_t cols,
                                                uint32_t tid, uint32_t nth, uint32_t off8 = 0) {
    const uint32_t cpr = cols >> 3;
    const bool pow2 = (cpr & (cpr - 1u)) == 0u;  // uniform: 16 / 32 / 64 / 128 columns take a shift instead of a division per chunk
    const uint32_t sh = 31u - (uint32_t)__clz((int)cpr);
    for (uint32_t q = tid; q < kTileRows * cpr; q += nth) {
        const uint32_t r = pow2 ? q >> sh : q / cpr, c8 = q - r * cpr, d8 = c8 + off8;
        const uint32_t grow = row0 + r;
        const bool ok = grow < B;
        const __half *src = g + (size_t)(ok ? grow : 0u) * cols + c8 * 8;
        const uint32_t dst = smem_u32(tiles + (d8 >> 3) * kActTileBytes + swz_off(r, d8 & 7u));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16u : 0u) : "memory");
    }
}

__device__ __forceinline__ uint32_t pack2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<const uint32_t *>(&h);
}

// ---- epilogues -----------------------------------------------------------------------------------------
// A CTA has 128 * NS threads: warp w owns TMEM lanes 32*(w%4).. (hardware rule) = batch rows, and column group
// w/4 of width CW = width/NS, so one accumulator row is drained by NS threads in parallel (the epilogue chain
// -- tcgen05.ld, ReLU, fp16 pack, swizzled st.shared -- is latency bound, not throughput bound).
__device__ __forceinline__ uint32_t relu2(float a, float b) {
    uint32_t r;  // one instruction: round both to fp16 with negative values clamped to +0 (low half = a)
    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
// The reference's other hidden activations (utils.h:425-470 forward, :538-580 backward from the stored POST-activation),
// evaluated on the fp32 accumulator; K_ACT = 10 (utils.h:41).  `act` is uniform over the launch.
constexpr float kKAct = 10.0f;
__device__ __forceinline__ float act_fwd(uint32_t act, float x) {
    switch (act) {
        case 1: return expf(x);
        case 2: return sinf(x);
        case 3: return 1.0f / (1.0f + expf(-x));
        case 4: { const float k = x * kKAct; return 0.5f * (k + sqrtf(k * k + 4.0f)) / kKAct; }
        case 5: return logf(expf(x * kKAct) + 1.0f) / kKAct;
        case 6: return x;
        default: return fmaxf(x, 0.0f);
    }
}
// d act / d pre-activation as a function of the post-activation y (what the tile holds)
__device__ __forceinline__ float act_bwd_from_out(uint32_t act, float y) {
    switch (act) {
        case 1: return y;
        case 3: return y * (1.0f - y);
        case 4: { const float k = y * kKAct; return k * k / (k * k + 1.0f); }
        case 5: return 1.0f - expf(-y * kKAct);
        case 6: return 1.0f;
        default: return y > 0.0f ? 1.0f : 0.0f;  // ReLU (Sine is rejected on the host: its derivative needs the pre-activation)
    }
}

// acc[row, col0 .. col0+CW) -> activation (ReLU: fast path) -> fp16 -> swizzled tile(s)
// generic activation: out of line, so that the ReLU kernels keep their register allocation and code size
template <int CW>
__device__ __noinline__ void epi_act_to_tile(uint32_t taddr, uint8_t *tiles, uint32_t row, uint32_t col0, uint32_t act) {
    float v[CW];
    tmem_ldN<CW>(taddr + col0, v);
    uint8_t *t = tiles + (col0 >> 6) * kActTileBytes + row * kTileRowBytes;
    const uint32_t c8 = (col0 & 63u) >> 3;
#pragma unroll
    for (int j = 0; j < CW; ++j) v[j] = act_fwd(act, v[j]);
#pragma unroll
    for (int j = 0; j < CW; j += 8) {
        uint4 u;
        u.x = pack2(v[j], v[j + 1]); u.y = pack2(v[j + 2], v[j + 3]); u.z = pack2(v[j + 4], v[j + 5]); u.w = pack2(v[j + 6], v[j + 7]);
        *reinterpret_cast<uint4 *>(t + (((c8 + (j >> 3)) ^ (row & 7u)) << 4)) = u;
    }
}
// to_tm: ALSO store the fp16 row piece into tensor memory at a_row + col0 / 2 (two K elements per 32-bit column, row = lane):
// the A operand of the next layer product when that product reads A from TMEM (CW == 32 only: one tcgen05.st.x16)
template <int CW>
__device__ __forceinline__ void epi_relu_to_tile(uint32_t taddr, uint8_t *tiles, uint32_t row, uint32_t col0, uint32_t act = 0u,
                                                 bool to_tm = false, uint32_t a_row = 0u) {
    if (act != 0u) { epi_act_to_tile<CW>(taddr, tiles, row, col0, act); return; }
    float v[CW];
    tmem_ldN<CW>(taddr + col0, v);
    uint8_t *t = tiles + (col0 >> 6) * kActTileBytes + row * kTileRowBytes;
    const uint32_t c8 = (col0 & 63u) >> 3;
    uint32_t p[CW / 2];
#pragma unroll
    for (int j = 0; j < CW / 2; ++j) p[j] = relu2(v[2 * j], v[2 * j + 1]);
#pragma unroll
    for (int j = 0; j < CW; j += 8)
        *reinterpret_cast<uint4 *>(t + (((c8 + (j >> 3)) ^ (row & 7u)) << 4)) = make_uint4(p[j >> 1], p[(j >> 1) + 1], p[(j >> 1) + 2], p[(j >> 1) + 3]);
    if constexpr (CW == 32) {
        if (to_tm) tmem_st16(a_row + (col0 >> 1), p);
    }
}

__device__ __forceinline__ uint32_t mask2(float a, float b, uint32_t h2) {
    // dL/dh * (h > 0): h is a ReLU output (>= 0, never -0), so "> 0" == "bit pattern != 0"
    // (one HSET2.BM + one LOP3 per pair; the bit-test form of the same predicate cost seven instructions)
    return pack2(a, b) & __hgt2_mask(*reinterpret_cast<const __half2 *>(&h2), __float2half2_rn(0.f));
}
// acc[row, col0 .. col0+CW) * (act[row, :] > 0) -> fp16 -> swizzled tile(s)   (ReLU backward from the saved activation)
__device__ __forceinline__ uint32_t dact2(uint32_t act, float a, float b, uint32_t h2) {
    const float2 y = __half22float2(*reinterpret_cast<const __half2 *>(&h2));
    return pack2(a * act_bwd_from_out(act, y.x), b * act_bwd_from_out(act, y.y));
}
template <int CW>
__device__ __noinline__ void epi_dact_to_tile(uint32_t taddr, const uint8_t *act_tiles, uint8_t *tiles, uint32_t row, uint32_t col0, uint32_t act) {
    float v[CW];
    tmem_ldN<CW>(taddr + col0, v);
    const uint32_t base = (col0 >> 6) * kActTileBytes + row * kTileRowBytes;
    const uint32_t c8 = (col0 & 63u) >> 3;
#pragma unroll
    for (int j = 0; j < CW; j += 8) {
        const uint32_t o = base + (((c8 + (j >> 3)) ^ (row & 7u)) << 4);
        const uint4 a = *reinterpret_cast<const uint4 *>(act_tiles + o);
        uint4 u;
        u.x = dact2(act, v[j], v[j + 1], a.x); u.y = dact2(act, v[j + 2], v[j + 3], a.y);
        u.z = dact2(act, v[j + 4], v[j + 5], a.z); u.w = dact2(act, v[j + 6], v[j + 7], a.w);
        *reinterpret_cast<uint4 *>(tiles + o) = u;
    }
}
template <int CW>
__device__ __forceinline__ void epi_mask_to_tile(uint32_t taddr, const uint8_t *act_tiles, uint8_t *tiles, uint32_t row, uint32_t col0,
                                                 uint32_t act = 0u, bool to_tm = false, uint32_t a_row = 0u) {
    if (act != 0u) { epi_dact_to_tile<CW>(taddr, act_tiles, tiles, row, col0, act); return; }
    float v[CW];
    tmem_ldN<CW>(taddr + col0, v);
    const uint32_t base = (col0 >> 6) * kActTileBytes + row * kTileRowBytes;
    const uint32_t c8 = (col0 & 63u) >> 3;
    uint32_t p[CW / 2];
#pragma unroll
    for (int j = 0; j < CW; j += 8) {
        const uint32_t o = base + (((c8 + (j >> 3)) ^ (row & 7u)) << 4);
        const uint4 a = *reinterpret_cast<const uint4 *>(act_tiles + o);
        uint4 u;
        u.x = mask2(v[j], v[j + 1], a.x); u.y = mask2(v[j + 2], v[j + 3], a.y);
        u.z = mask2(v[j + 4], v[j + 5], a.z); u.w = mask2(v[j + 6], v[j + 7], a.w);
        *reinterpret_cast<uint4 *>(tiles + o) = u;
        p[j >> 1] = u.x; p[(j >> 1) + 1] = u.y; p[(j >> 1) + 2] = u.z; p[(j >> 1) + 3] = u.w;
    }
    if constexpr (CW == 32) {
        if (to_tm) tmem_st16(a_row + (col0 >> 1), p);
    }
}
// acc[row, 0..n) -> fp16 -> global row (n % 16 == 0; 32 B per 16 columns, consecutive rows are contiguous)
__device__ __forceinline__ void epi_to_global(uint32_t taddr, __half *__restrict__ dst_row, uint32_t cbeg, uint32_t cend, bool write) {
    for (uint32_t c0 = cbeg; c0 < cend; c0 += 16) {
        float v[16];
        tmem_ld16(taddr + c0, v);
        if (write) {
            uint4 lo, hi;
            lo.x = pack2(v[0], v[1]); lo.y = pack2(v[2], v[3]); lo.z = pack2(v[4], v[5]); lo.w = pack2(v[6], v[7]);
            hi.x = pack2(v[8], v[9]); hi.y = pack2(v[10], v[11]); hi.z = pack2(v[12], v[13]); hi.w = pack2(v[14], v[15]);
            reinterpret_cast<uint4 *>(dst_row + c0)[0] = lo;
            reinterpret_cast<uint4 *>(dst_row + c0)[1] = hi;
        }
    }
}

}  // namespace

// hidden_dim 256 (ffmlp_wide.cu): weights streamed through shared memory, two-phase backward
int wide_forward(const char *fn, const void *inputs, const void *weights, int wdtype, uint32_t B, uint32_t in, uint32_t out, uint32_t nl,
                 uint32_t act, void *forward_buffer, void *outputs, cudaStream_t st);
int wide_backward(const char *fn, const void *grad, const void *inputs, const void *weights, int wdtype, const void *forward_buffer, uint32_t B,
                  uint32_t in, uint32_t out, uint32_t nl, uint32_t act, int calc_dx, void *backward_buffer, void *grad_inputs, void *grad_weights,
                  int gw_f32, float *ws, cudaStream_t st);
}  // namespace ngp
