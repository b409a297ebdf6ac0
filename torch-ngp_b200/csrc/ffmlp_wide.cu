// ffmlp_wide.cu -- FFMLP with hidden_dim = 256 on tcgen05 (sm_100a): the width whose weights do not fit in shared memory.
//
// Contract: /root/reference/ffmlp/src/ffmlp.cu, the WIDTH = 256 instantiations (forward :657, inference :695, backward :834)
// and the host flow of ffmlp_backward (:749-895): a fused kernel produces dL/d(pre-activation) of every hidden layer into
// `backward_buffer`, then one batch-wide product per matrix gives the weight gradients from `forward_buffer` /
// `backward_buffer` (the reference: CUTLASS split-k GEMMs on a stream pool, fp16 output).  Same two-phase structure here,
// because at this width neither the weights (a 256 x 256 fp16 matrix is 128 KB) nor the weight-gradient accumulators (one
// 256 x 256 fp32 matrix is all 512 TMEM columns) can stay on the SM next to a batch tile:
//   k_ffmlp_wide_fwd   one 128-row tile per CTA at a time; every weight matrix is STREAMED through two 32 KB shared-memory
//                      buffers in 64-wide K chunks ([N rows x 64] K-major images) while the tensor core works on the
//                      previous chunk; D[128 x 256] fp32 in TMEM; the drain (tcgen05.ld -> activation -> fp16) writes the
//                      next layer's A operand into the other activation buffer and, in training mode, the row into
//                      forward_buffer[l] (the reference's layout: [num_layers, B, 256]).
//   k_ffmlp_wide_bwd   dY tile -> dH_{nl-1} = dY W_last -> G_l = dH_l * act'(H_l) (H_l from forward_buffer) -> dH_{l-1} = G_l W_l
//                      ... -> dX; the weight rows of a K chunk are staged as MN-major B images ([64 k-rows x N]); every G_l
//                      goes to backward_buffer[l].
//   k_ffmlp_wide_dw    dW = G^T A over the whole batch, split over the CTAs along the batch: 64-row stages of both operands
//                      through a 3-deep cp.async ring, M = 128 x N <= 256 MN-major products accumulating a full
//                      [256 x 256] fp32 matrix in TMEM (512 columns), one vector-reduction flush per CTA; the last CTA
//                      converts the sum to the requested dtype and leaves the workspace zeroed.
#include "ffmlp_tiles.cuh"

namespace ngp {
namespace {

using namespace umma;

constexpr uint32_t kWideActBytes = 4 * kActTileBytes;  // 128 rows x 256 halves
constexpr uint32_t kChunkBytes = 32 * 1024;            // [256 x 64] halves
constexpr uint32_t kSlabBytes = 64 * kTileRowBytes;    // 64 k-rows x 64 halves of an MN-major image

template <typename WT> __device__ __forceinline__ uint4 load8(const WT *p) {
    if constexpr (sizeof(WT) == 2) {
        return __ldg(reinterpret_cast<const uint4 *>(p));
    } else {
        const float4 a = __ldg(reinterpret_cast<const float4 *>(p)), b = __ldg(reinterpret_cast<const float4 *>(p) + 1);
        uint4 v;
        v.x = pack2(a.x, a.y); v.y = pack2(a.z, a.w); v.z = pack2(b.x, b.y); v.w = pack2(b.z, b.w);
        return v;
    }
}
// rows [0, N) of W (row-major [N x ld]), columns [k0, k0 + kc) -> K-major swizzled image [N x 64]
template <typename WT>
__device__ __forceinline__ void load_chunk_kmajor(uint8_t *dst, const WT *__restrict__ W, uint32_t N, uint32_t ld, uint32_t k0, uint32_t kc,
                                                  uint32_t tid, uint32_t nth) {
    const uint32_t cpr = kc >> 3;
    for (uint32_t q = tid; q < N * cpr; q += nth) {
        const uint32_t r = q / cpr, c8 = q - r * cpr;
        *reinterpret_cast<uint4 *>(dst + swz_off(r, c8)) = load8<WT>(W + (size_t)r * ld + k0 + c8 * 8);
    }
}
// rows [k0, k0 + kc) of W (row-major [. x N]), all N columns -> MN-major image: N/64 slabs of [kc x 64]
template <typename WT>
__device__ __forceinline__ void load_chunk_mnmajor(uint8_t *dst, const WT *__restrict__ W, uint32_t N, uint32_t k0, uint32_t kc, uint32_t tid,
                                                   uint32_t nth) {
    const uint32_t cpr = N >> 3;
    for (uint32_t q = tid; q < kc * cpr; q += nth) {
        const uint32_t r = q / cpr, c8 = q - r * cpr;
        *reinterpret_cast<uint4 *>(dst + (c8 >> 3) * kSlabBytes + swz_off(r, c8 & 7u)) = load8<WT>(W + (size_t)(k0 + r) * N + c8 * 8);
    }
}

// the two weight buffers: a commit on s_free[b] follows every group of MMAs that read buffer b; exactly one wait consumes it
struct ring_t {
    uint32_t ph[2] = {0u, 0u};
    bool pend[2] = {false, false};
    __device__ __forceinline__ void wait(uint64_t *bars, uint32_t b) {
        if (pend[b]) {
            mbar_wait(&bars[b], ph[b]);
            ph[b] ^= 1u;
            pend[b] = false;
        }
    }
};

// acc[row, c0 .. c0+32) -> activation -> fp16 -> swizzled tile (+ the global row of forward_buffer)
template <bool GACT>
__device__ __forceinline__ void epi_hidden32(uint32_t taddr, uint8_t *tiles, uint32_t row, uint32_t c0, uint32_t act, __half *grow) {
    float v[32];
    tmem_ld32(taddr + c0, v);
    uint32_t p[16];
    if (GACT && act != 0u) {
#pragma unroll
        for (int j = 0; j < 32; j += 2) p[j >> 1] = pack2(act_fwd(act, v[j]), act_fwd(act, v[j + 1]));
    } else {
#pragma unroll
        for (int j = 0; j < 32; j += 2) p[j >> 1] = relu2(v[j], v[j + 1]);
    }
    uint8_t *t = tiles + (c0 >> 6) * kActTileBytes + row * kTileRowBytes;
    const uint32_t c8 = (c0 & 63u) >> 3;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint4 u = make_uint4(p[4 * q], p[4 * q + 1], p[4 * q + 2], p[4 * q + 3]);
        *reinterpret_cast<uint4 *>(t + (((c8 + q) ^ (row & 7u)) << 4)) = u;
        if (grow) *reinterpret_cast<uint4 *>(grow + c0 + 8 * q) = u;
    }
}

// ==========================================================================================================
// forward / inference
// ==========================================================================================================
template <typename WT, bool GACT>
__global__ void __launch_bounds__(256) k_ffmlp_wide_fwd(const __half *__restrict__ x, const WT *__restrict__ w, __half *__restrict__ y,
                                                        __half *__restrict__ fbuf, const mlp_t m, uint32_t ntiles, const int *__restrict__ limit) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_free[2];
    const uint32_t Beff = eff_batch(m.B, limit);
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t row = (warp & 3u) * 32u + lane, ch = warp >> 2;
    uint8_t *act[2] = {smem, smem + kWideActBytes};
    uint8_t *wb[2] = {smem + 2 * kWideActBytes, smem + 2 * kWideActBytes + kChunkBytes};
    if (warp == 0) tmem_alloc(&s_tmem, 256);
    if (tid < 2) mbar_init(&s_free[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t trow = tmem + (((warp & 3u) * 32u) << 16);
    ring_t ring;
    uint32_t jc = 0;  // running chunk counter: chunk j lives in weight buffer j & 1

    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const uint32_t grow = tile * kTileRows + row;
        load_rows_async(act[0], x, tile * kTileRows, Beff, m.in, tid, 256);
        cp_async_commit();
        ring.wait(s_free, jc & 1u);
        load_chunk_kmajor<WT>(wb[jc & 1u], w, m.rows[0], m.K[0], 0, min(64u, m.K[0]), tid, 256);
        uint32_t cur = 0;
        for (uint32_t l = 0; l <= m.nl; ++l) {
            const uint32_t N = m.rows[l], K = m.K[l], nk = ktiles(K);
            const uint32_t idesc = make_idesc(128, N, false, false);
            for (uint32_t kt = 0; kt < nk; ++kt, ++jc) {
                const uint32_t b = jc & 1u, kc = min(64u, K - kt * 64u);
                cp_async_wait<0>();
                fence_proxy_async();
                fence_before_sync();
                __syncthreads();
                if (tid == 0) {
                    fence_after_sync();
                    const uint32_t a_base = smem_u32(act[cur]) + kt * kActTileBytes, b_base = smem_u32(wb[b]);
                    for (uint32_t kk = 0; kk < kc / 16; ++kk)
                        mma_f16(tmem, make_desc(a_base + kk * 32u, 0, kAtomBytes), make_desc(b_base + kk * 32u, 0, kAtomBytes), idesc, (kt | kk) ? 1u : 0u);
                    mma_commit(&s_free[b]);
                }
                ring.pend[b] = true;
                // stage the next chunk (same layer, or the first of the next layer) while the tensor core works
                uint32_t nl_ = l, nkt = kt + 1;
                if (nkt == nk) { nl_ = l + 1; nkt = 0; }
                if (nl_ <= m.nl) {
                    ring.wait(s_free, b ^ 1u);
                    load_chunk_kmajor<WT>(wb[b ^ 1u], w + m.woff[nl_], m.rows[nl_], m.K[nl_], nkt * 64u, min(64u, m.K[nl_] - nkt * 64u), tid, 256);
                }
            }
            ring.wait(s_free, (jc - 1u) & 1u);  // the layer's accumulator is complete
            fence_after_sync();
            if (l < m.nl) {
                __half *g = (fbuf && grow < Beff) ? fbuf + ((size_t)l * m.B + grow) * 256u : nullptr;
#pragma unroll 1
                for (uint32_t c0 = ch * 128u; c0 < ch * 128u + 128u; c0 += 32u) epi_hidden32<GACT>(trow, act[cur ^ 1u], row, c0, m.act, g);
                cur ^= 1u;
            } else {
                for (uint32_t c0 = ch * 16u; c0 < m.out; c0 += 32u) epi_to_global(trow, y + (size_t)grow * m.out, c0, c0 + 16u, grow < Beff);
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ==========================================================================================================
// backward, phase 1: dL/d(pre-activation) of every hidden layer (-> backward_buffer) and dL/dx
// ==========================================================================================================
template <typename WT, bool GACT>
__global__ void __launch_bounds__(256) k_ffmlp_wide_bwd(const __half *__restrict__ grad, const WT *__restrict__ w, const __half *__restrict__ fbuf,
                                                        __half *__restrict__ bbuf, __half *__restrict__ grad_x, const mlp_t m, uint32_t ntiles,
                                                        int calc_dx, const int *__restrict__ limit) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_free[2];
    const uint32_t Beff = eff_batch(m.B, limit);
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t row = (warp & 3u) * 32u + lane, ch = warp >> 2;
    uint8_t *act[2] = {smem, smem + kWideActBytes};
    uint8_t *wb[2] = {smem + 2 * kWideActBytes, smem + 2 * kWideActBytes + kChunkBytes};
    if (warp == 0) tmem_alloc(&s_tmem, 256);
    if (tid < 2) mbar_init(&s_free[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t trow = tmem + (((warp & 3u) * 32u) << 16);
    ring_t ring;
    uint32_t jc = 0;
    const int s_end = calc_dx ? 0 : 1;  // the last product (dX = G_0 W_0) only when asked for

    for (uint32_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const uint32_t grow = tile * kTileRows + row;
        const bool live = grow < Beff;
        load_rows_async(act[0], grad, tile * kTileRows, Beff, m.out, tid, 256);
        cp_async_commit();
        ring.wait(s_free, jc & 1u);
        load_chunk_mnmajor<WT>(wb[jc & 1u], w + m.woff[m.nl], m.K[m.nl], 0, min(64u, m.rows[m.nl]), tid, 256);
        uint32_t cur = 0;
        for (int s = (int)m.nl; s >= s_end; --s) {  // product with W_s: dH_{s-1} [128 x K_s] = G_s [128 x rows_s] W_s [rows_s x K_s]
            const uint32_t N = m.K[s], K = m.rows[s], nk = ktiles(K);
            const uint32_t idesc = make_idesc(128, N, false, true);
            const uint32_t b_lbo = N > 64u ? kSlabBytes : 0u;
            // H_{s-1}[row, this thread's 128 columns]: requested now, used in the drain below
            uint4 hreg[16];
            if (s > 0) {
                const uint4 *hp = reinterpret_cast<const uint4 *>(fbuf + ((size_t)(s - 1) * m.B + (live ? grow : 0u)) * 256u + ch * 128u);
#pragma unroll
                for (int q = 0; q < 16; ++q) hreg[q] = live ? __ldg(hp + q) : make_uint4(0u, 0u, 0u, 0u);
            }
            for (uint32_t kt = 0; kt < nk; ++kt, ++jc) {
                const uint32_t b = jc & 1u, kc = min(64u, K - kt * 64u);
                cp_async_wait<0>();
                fence_proxy_async();
                fence_before_sync();
                __syncthreads();
                if (tid == 0) {
                    fence_after_sync();
                    const uint32_t a_base = smem_u32(act[cur]) + kt * kActTileBytes, b_base = smem_u32(wb[b]);
                    for (uint32_t kk = 0; kk < kc / 16; ++kk)
                        mma_f16(tmem, make_desc(a_base + kk * 32u, 0, kAtomBytes), make_desc(b_base + kk * 2048u, b_lbo, kAtomBytes), idesc, (kt | kk) ? 1u : 0u);
                    mma_commit(&s_free[b]);
                }
                ring.pend[b] = true;
                int ns = s;
                uint32_t nkt = kt + 1;
                if (nkt == nk) { ns = s - 1; nkt = 0; }
                if (ns >= s_end) {
                    ring.wait(s_free, b ^ 1u);
                    load_chunk_mnmajor<WT>(wb[b ^ 1u], w + m.woff[ns], m.K[ns], nkt * 64u, min(64u, m.rows[ns] - nkt * 64u), tid, 256);
                }
            }
            ring.wait(s_free, (jc - 1u) & 1u);
            fence_after_sync();
            if (s > 0) {  // G_{s-1} = dH_{s-1} * act'(H_{s-1}) -> next A operand + backward_buffer[s-1]
                __half *g = live ? bbuf + ((size_t)(s - 1) * m.B + grow) * 256u : nullptr;
                uint8_t *tiles = act[cur ^ 1u];
#pragma unroll
                for (int grp = 0; grp < 4; ++grp) {
                    const uint32_t c0 = ch * 128u + grp * 32u;
                    float v[32];
                    tmem_ld32(trow + c0, v);
                    uint8_t *t = tiles + (c0 >> 6) * kActTileBytes + row * kTileRowBytes;
                    const uint32_t c8 = (c0 & 63u) >> 3;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const uint4 h = hreg[grp * 4 + q];
                        uint4 u;
                        if (GACT && m.act != 0u) {
                            u.x = dact2(m.act, v[8 * q], v[8 * q + 1], h.x); u.y = dact2(m.act, v[8 * q + 2], v[8 * q + 3], h.y);
                            u.z = dact2(m.act, v[8 * q + 4], v[8 * q + 5], h.z); u.w = dact2(m.act, v[8 * q + 6], v[8 * q + 7], h.w);
                        } else {
                            u.x = mask2(v[8 * q], v[8 * q + 1], h.x); u.y = mask2(v[8 * q + 2], v[8 * q + 3], h.y);
                            u.z = mask2(v[8 * q + 4], v[8 * q + 5], h.z); u.w = mask2(v[8 * q + 6], v[8 * q + 7], h.w);
                        }
                        *reinterpret_cast<uint4 *>(t + (((c8 + q) ^ (row & 7u)) << 4)) = u;
                        if (g) *reinterpret_cast<uint4 *>(g + c0 + 8 * q) = u;
                    }
                }
                cur ^= 1u;
            } else {
                for (uint32_t c0 = ch * 16u; c0 < m.in; c0 += 32u) epi_to_global(trow, grad_x + (size_t)grow * m.in, c0, c0 + 16u, live);
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}

// ==========================================================================================================
// backward, phase 2: dW [Md x Nd] = P^T Q over the batch (P = G_i [B x Md], Q = layer input [B x Nd])
// ==========================================================================================================
constexpr uint32_t kDwStageRows = 64, kDwStages = 3, kDwStageBytes = 2 * kChunkBytes;

__device__ __forceinline__ void dw_load_stage(uint8_t *slot, const __half *__restrict__ P, uint32_t Md, const __half *__restrict__ Q, uint32_t Nd,
                                              uint32_t r0, uint32_t Beff, uint32_t tid, uint32_t nth) {
    const uint32_t cp = Md >> 3, cq = Nd >> 3;
    for (uint32_t q = tid; q < kDwStageRows * cp; q += nth) {
        const uint32_t r = q / cp, c8 = q - r * cp;
        const bool ok = r0 + r < Beff;
        const __half *src = P + (size_t)(ok ? r0 + r : 0u) * Md + c8 * 8;
        const uint32_t dst = smem_u32(slot + (c8 >> 3) * kSlabBytes + swz_off(r, c8 & 7u));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16u : 0u) : "memory");
    }
    for (uint32_t q = tid; q < kDwStageRows * cq; q += nth) {
        const uint32_t r = q / cq, c8 = q - r * cq;
        const bool ok = r0 + r < Beff;
        const __half *src = Q + (size_t)(ok ? r0 + r : 0u) * Nd + c8 * 8;
        const uint32_t dst = smem_u32(slot + kChunkBytes + (c8 >> 3) * kSlabBytes + swz_off(r, c8 & 7u));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16u : 0u) : "memory");
    }
}

__global__ void __launch_bounds__(256) k_ffmlp_wide_dw(const __half *__restrict__ P, uint32_t Md, const __half *__restrict__ Q, uint32_t Nd,
                                                       float *__restrict__ ws_mat, void *__restrict__ gw_mat, int gw_f32, uint32_t *__restrict__ counter,
                                                       uint32_t B, uint32_t tmem_cols, const int *__restrict__ limit) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem, s_last;
    __shared__ __align__(8) uint64_t s_free[kDwStages];
    const uint32_t Beff = eff_batch(B, limit);
    const uint32_t nst = div_round_up(Beff, kDwStageRows);
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t row = (warp & 3u) * 32u + lane, ch = warp >> 2;
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid < kDwStages) mbar_init(&s_free[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t trow = tmem + (((warp & 3u) * 32u) << 16);
    const uint32_t mine = blockIdx.x < nst ? (nst - blockIdx.x + gridDim.x - 1) / gridDim.x : 0u;  // stages of this CTA
    const uint32_t nh = Md > 128u ? 2u : 1u;
    const uint32_t idesc = make_idesc(128, Nd, true, true);
    const uint32_t a_lbo = Md > 64u ? kSlabBytes : 0u, b_lbo = Nd > 64u ? kSlabBytes : 0u;
    uint32_t ph[kDwStages] = {0u, 0u, 0u};

    for (uint32_t i = 0; i < kDwStages - 1; ++i) {
        if (i < mine) dw_load_stage(smem + i * kDwStageBytes, P, Md, Q, Nd, (blockIdx.x + i * gridDim.x) * kDwStageRows, Beff, tid, 256);
        cp_async_commit();
    }
    for (uint32_t it = 0; it < mine; ++it) {
        const uint32_t slot = it % kDwStages;
        cp_async_wait<kDwStages - 2>();
        fence_proxy_async();
        fence_before_sync();
        __syncthreads();
        if (tid == 0) {
            fence_after_sync();
            const uint32_t p_base = smem_u32(smem + slot * kDwStageBytes), q_base = p_base + kChunkBytes;
            for (uint32_t kk = 0; kk < kDwStageRows / 16; ++kk)
                for (uint32_t h = 0; h < nh; ++h)
                    mma_f16(tmem + h * Nd, make_desc(p_base + h * 2u * kSlabBytes + kk * 2048u, a_lbo, kAtomBytes),
                            make_desc(q_base + kk * 2048u, b_lbo, kAtomBytes), idesc, (it | kk) ? 1u : 0u);
            mma_commit(&s_free[slot]);
        }
        // refill the slot the PREVIOUS iteration's products read
        const uint32_t nxt = it + kDwStages - 1;
        if (nxt < mine) {
            const uint32_t ns = nxt % kDwStages;
            if (it >= 1) {
                mbar_wait(&s_free[ns], ph[ns]);
                ph[ns] ^= 1u;
            }
            dw_load_stage(smem + ns * kDwStageBytes, P, Md, Q, Nd, (blockIdx.x + nxt * gridDim.x) * kDwStageRows, Beff, tid, 256);
        }
        cp_async_commit();
    }
    cp_async_wait<0>();
    if (mine) {
        // the last commit covers every product of this CTA; earlier, unconsumed phases of the other slots are left behind
        const uint32_t ls = (mine - 1) % kDwStages;
        // phases consumed on ls so far: one per refill of that slot
        mbar_wait(&s_free[ls], ph[ls]);
        fence_after_sync();
        for (uint32_t h = 0; h < nh; ++h) {
            const uint32_t orow = h * 128u + row;
            for (uint32_t c0 = ch * 16u; c0 < Nd; c0 += 32u) {  // warp-uniform trip count: tcgen05.ld is .sync.aligned
                float v[16];
                tmem_ld16(trow + h * Nd + c0, v);
                if (orow < Md) {
                    float *d = ws_mat + (size_t)orow * Nd + c0;
#pragma unroll
                    for (int j = 0; j < 16; j += 4)
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]), "f"(v[j + 3]) : "memory");
                }
            }
        }
    }
    fence_before_sync();
    __threadfence();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, tmem_cols);
    if (tid == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (uint32_t i = tid; i < Md * Nd; i += blockDim.x) {
        const float v = __ldcg(ws_mat + i);
        if (gw_f32) reinterpret_cast<float *>(gw_mat)[i] = v;
        else reinterpret_cast<__half *>(gw_mat)[i] = __float2half_rn(v);
        ws_mat[i] = 0.f;
    }
    if (tid == 0) *counter = 0u;
}

template <typename K> int set_smem_wide(K kernel, size_t bytes) {
    NGP_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return NGP_OK;
}

int check_wide(const char *fn, uint32_t in, uint32_t out, uint32_t nl) {
    NGP_CHECK_ARG(in > 0 && in % 16 == 0 && in <= 256, "%s: input_dim should be 16 * m, 0 < m <= 16 at hidden_dim 256, but got %u", fn, in);
    NGP_CHECK_ARG(out > 0 && out % 16 == 0 && out <= 128, "%s: (padded) output_dim should be 16 * m, 0 < m <= 8, but got %u", fn, out);
    NGP_CHECK_ARG(nl >= 2 && nl <= kMaxLayers, "%s: num_layers should be in [2, %u], but got %u", fn, kMaxLayers, nl);
    return NGP_OK;
}

}  // namespace

int wide_forward(const char *fn, const void *inputs, const void *weights, int wdtype, uint32_t B, uint32_t in, uint32_t out, uint32_t nl,
                 uint32_t act, void *forward_buffer, void *outputs, cudaStream_t st) {
    int rc = check_wide(fn, in, out, nl);
    if (rc != NGP_OK) return rc;
    mlp_t m;
    make_mlp(m, B, in, out, 256, nl);
    m.act = act;
    const uint32_t ntiles = div_round_up(B, kTileRows);
    const uint32_t grid = min(ntiles, (uint32_t)sm_count());
    const size_t smem = 2 * kWideActBytes + 2 * kChunkBytes;
#define NGP_WFWD(WT_, G_)                                                                                                         \
    {                                                                                                                             \
        rc = set_smem_wide(k_ffmlp_wide_fwd<WT_, G_>, smem);                                                                      \
        if (rc != NGP_OK) return rc;                                                                                              \
        NGP_PROF("k_ffmlp_wide_fwd", st);                                                                                         \
        k_ffmlp_wide_fwd<WT_, G_><<<grid, 256, smem, st>>>((const __half *)inputs, (const WT_ *)weights, (__half *)outputs,       \
                                                          (__half *)forward_buffer, m, ntiles, batch_limit());                   \
    }
    if (wdtype == NGP_DTYPE_F16) { if (act) NGP_WFWD(__half, true) else NGP_WFWD(__half, false) }
    else { if (act) NGP_WFWD(float, true) else NGP_WFWD(float, false) }
#undef NGP_WFWD
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int wide_backward(const char *fn, const void *grad, const void *inputs, const void *weights, int wdtype, const void *forward_buffer, uint32_t B,
                  uint32_t in, uint32_t out, uint32_t nl, uint32_t act, int calc_dx, void *backward_buffer, void *grad_inputs, void *grad_weights,
                  int gw_f32, float *ws, cudaStream_t st) {
    int rc = check_wide(fn, in, out, nl);
    if (rc != NGP_OK) return rc;
    NGP_CHECK_ARG(forward_buffer && backward_buffer,
                  "%s: hidden_dim 256 needs forward_buffer (written by ffmlp_forward) and backward_buffer, [num_layers, B, 256] half each, as the "
                  "reference allocates them (ffmlp.py:33, :70): at this width the activations cannot stay on the SM", fn);
    mlp_t m;
    make_mlp(m, B, in, out, 256, nl);
    m.act = act;
    const uint32_t ntiles = div_round_up(B, kTileRows);
    const uint32_t grid = min(ntiles, (uint32_t)sm_count());
    const size_t smem = 2 * kWideActBytes + 2 * kChunkBytes;
#define NGP_WBWD(WT_, G_)                                                                                                         \
    {                                                                                                                             \
        rc = set_smem_wide(k_ffmlp_wide_bwd<WT_, G_>, smem);                                                                      \
        if (rc != NGP_OK) return rc;                                                                                              \
        NGP_PROF("k_ffmlp_wide_bwd", st);                                                                                         \
        k_ffmlp_wide_bwd<WT_, G_><<<grid, 256, smem, st>>>((const __half *)grad, (const WT_ *)weights, (const __half *)forward_buffer, \
                                                          (__half *)backward_buffer, (__half *)grad_inputs, m, ntiles, calc_dx, batch_limit()); \
    }
    if (wdtype == NGP_DTYPE_F16) { if (act) NGP_WBWD(__half, true) else NGP_WBWD(__half, false) }
    else { if (act) NGP_WBWD(float, true) else NGP_WBWD(float, false) }
#undef NGP_WBWD
    NGP_CHECK_LAUNCH();
    // ---- weight gradients, one batch-wide product per matrix ----
    rc = set_smem_wide(k_ffmlp_wide_dw, kDwStages * kDwStageBytes);
    if (rc != NGP_OK) return rc;
    uint32_t *counter = reinterpret_cast<uint32_t *>(ws + ((m.nweights + 3u) & ~3u));
    const uint32_t nst = div_round_up(B, kDwStageRows);
    const uint32_t gdw = min(nst, (uint32_t)sm_count());
    const size_t esz = gw_f32 ? 4 : 2;
    for (uint32_t i = 0; i <= nl; ++i) {
        const __half *P = i == nl ? (const __half *)grad : (const __half *)backward_buffer + (size_t)i * B * 256u;
        const __half *Q = i == 0 ? (const __half *)inputs : (const __half *)forward_buffer + (size_t)(i - 1) * B * 256u;
        const uint32_t Md = m.rows[i], Nd = m.K[i];
        const uint32_t cols = tmem_cols_pow2((Md > 128u ? 2u : 1u) * Nd);
        NGP_PROF("k_ffmlp_wide_dw", st);
        k_ffmlp_wide_dw<<<gdw, 256, kDwStages * kDwStageBytes, st>>>(P, Md, Q, Nd, ws + m.woff[i], (uint8_t *)grad_weights + (size_t)m.woff[i] * esz, gw_f32,
                                                                      counter, B, cols, batch_limit());
    }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

}  // namespace ngp
