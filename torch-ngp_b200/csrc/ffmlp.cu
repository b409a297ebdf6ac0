// ffmlp.cu -- fully-fused small MLP on tcgen05 tensor cores (sm_100a).
//
// Contract: /root/reference/ffmlp/src/ffmlp.cu (kernel_mlp_fused :331-407, kernel_mlp_fused_backward :410-518,
// host ffmlp_forward :635-671, ffmlp_inference :673-709, ffmlp_backward :749-895) behind the C ABI of
// include/ngp_b200.h:  h = ReLU(W_0 x), h = ReLU(W_k h) x (num_layers-1), y = W_last h, no biases, fp16 storage.
//
// Design (not a translation of the reference's wmma register-fragment kernel):
//   * one CTA = 128 threads = one 128-row batch tile at a time, persistent over tiles; thread t owns row t
//     (= TMEM lane t) in every epilogue;
//   * every weight matrix is staged ONCE per CTA into shared memory as 128B-swizzled UMMA tiles
//     (fp32 parameters are rounded to fp16 on the way in: no per-call cast kernel);
//   * each layer is a few tcgen05.mma (M=128, N=width, K=16) issued by one thread, accumulating in fp32 in
//     TMEM; the epilogue (tcgen05.ld -> ReLU -> fp16) writes the next layer's A operand straight back into a
//     swizzled shared-memory tile -- activations never touch HBM (the reference streams
//     [num_layers, B, hidden] out in training mode and reads it back in the backward);
//   * the backward recomputes the forward per tile, keeps every activation / gradient tile in shared
//     memory, runs dL/dx through MN-major views of the SAME weight images, and accumulates all weight
//     gradients dW_m += G_m^T H_{m-1} in TMEM across the CTA's tiles (fp32, K = whole batch share), flushed
//     once per CTA with vector reductions; the last CTA converts the fp32 sum to the requested dtype.
//   * input / upstream-gradient tiles are prefetched one tile ahead with cp.async (zero-filled ragged tail).
#include "ffmlp_tiles.cuh"

namespace ngp {
namespace {

using namespace umma;

// ---- MMA issue helpers (one thread) ----------------------------------------------------------------------
// D[128 x N] = A[128 x K] (K-major activation tiles) * W^T, W = weight matrix image with N rows (K-major B)
__device__ __forceinline__ void issue_fwd(uint32_t d_tmem, const uint8_t *a_tiles, const uint8_t *w_img, uint32_t N, uint32_t K) {
    issue_kmajor(d_tmem, smem_u32(a_tiles), kActTileBytes, smem_u32(w_img), N * kTileRowBytes, N, K, 0);
}
// D[128 x N] = G[128 x K] (K-major tiles) * W, W = weight image [K rows x N cols] read as an MN-major B operand
// (N <= 64: one slab; N = 128: two slabs, LBO = slab stride = K * 128 bytes)
__device__ __forceinline__ void issue_bwd_dx(uint32_t d_tmem, const uint8_t *g_tiles, const uint8_t *w_img, uint32_t N, uint32_t K) {
    const uint32_t idesc = make_idesc(128, N, false, true);
    const uint32_t a_base = smem_u32(g_tiles), b_base = smem_u32(w_img);
    const uint32_t b_lbo = N > 64 ? K * kTileRowBytes : 0u;
    for (uint32_t k16 = 0; k16 < K / 16; ++k16) {
        const uint64_t ad = make_desc(a_base + (k16 >> 2) * kActTileBytes + (k16 & 3u) * 32u, 0, kAtomBytes);
        const uint64_t bd = make_desc(b_base + k16 * 2048u, b_lbo, kAtomBytes);
        mma_f16(d_tmem, ad, bd, idesc, k16 > 0 ? 1u : 0u);
    }
}
// D[M x N] (+)= G^T H over the tile's 128 rows: G tiles [128 x M] as MN-major A (M <= 64: the second 64-wide
// slab aliases the first, rows >= 64 of D are duplicates and ignored), H tiles [128 x N] as MN-major B
__device__ __forceinline__ void issue_dw(uint32_t d_tmem, const uint8_t *g_tiles, const uint8_t *h_tiles, uint32_t M, uint32_t N, uint32_t acc) {
    const uint32_t idesc = make_idesc(128, N, true, true);
    const uint32_t a_base = smem_u32(g_tiles), b_base = smem_u32(h_tiles);
    const uint32_t a_lbo = M > 64 ? kActTileBytes : 0u, b_lbo = N > 64 ? kActTileBytes : 0u;
    for (uint32_t k16 = 0; k16 < kTileRows / 16; ++k16) {
        const uint64_t ad = make_desc(a_base + k16 * 2048u, a_lbo, kAtomBytes);
        const uint64_t bd = make_desc(b_base + k16 * 2048u, b_lbo, kAtomBytes);
        mma_f16(d_tmem, ad, bd, idesc, (k16 > 0 || acc) ? 1u : 0u);
    }
}


// ==========================================================================================================
// network_ff glue fused into the MLP kernels (/root/reference/nerf/network_ff.py:51-74, activation.py:5-17)
// ==========================================================================================================
// IO = 0  plain FFMLP (the reference operator).
// IO = 1  sigma net forward: besides h [M,16] the drain writes sigmas = density_scale * exp(h0)  (trunc_exp, network_ff.py:60).
// IO = 2  colour net: the A operand tile [SH4(dir) | h1..h15 | 0] (network_ff.py:63-67, shencoder.cu:27-120) is BUILT in shared
//         memory from dirs [M,3] f32 and h [M,16] f16 -- the 64 B/sample `cin` matrix never exists in HBM; the forward drain
//         writes rgbs = sigmoid(c0..c2) (rounded to half like torch.sigmoid of a half tensor, network_ff.py:73) instead of c;
//         the backward builds its upstream tile from grad_rgbs * rgb (1 - rgb) and its dL/dx drain writes
//         grad_h = [grad_sigma * density_scale * exp(clamp(h0, -15, 15)) | dL/dcin[16..30]] (activation.py:12-17) -- so the four
//         elementwise kernels between the two nets (k_nerf_mid_fwd/bwd, k_nerf_rgb_fwd/bwd) and c / grad_c / grad_cin are gone.
// Arithmetic is that of those kernels, expression for expression (nerf_step.cu): results are bit-identical.
struct nerf_io_t {
    const float *dirs;         // [M,3]
    float *sigmas;             // IO 1 out
    float *rgbs;               // IO 2 forward out; IO 2 backward in
    const float *grad_rgbs;    // IO 2 backward in [M,3]
    const float *grad_sigmas;  // IO 2 backward in [M]
    float density_scale;
    const float *target;       // IO 3 (SDF regression): [B] f32
    const float *loss_scale;   // IO 3: device scalar or nullptr
    float *loss_out;           // IO 3: device scalar, accumulated into
};
__device__ __forceinline__ float half_round(float v) { return __half2float(__float2half_rn(v)); }

// row `row` of the colour net's input tile.  parts bit 0: SH chunks (columns 0..15), bit 1: geo chunks (columns 16..31).
// Returns the word holding (h0, h1) when the geo part was built (the backward's trunc_exp gradient needs h0).
__device__ __forceinline__ uint32_t nerf_color_row(uint8_t *tile, const float *__restrict__ dirs, const __half *__restrict__ h, uint32_t grow,
                                                   bool ok, uint32_t row, uint32_t parts) {
    uint8_t *t = tile + row * kTileRowBytes;
    const uint32_t sw = row & 7u;
    uint32_t h01 = 0u;
    if (parts & 1u) {
        float x = 0.f, y = 0.f, z = 0.f;
        if (ok) { x = dirs[3 * (size_t)grow]; y = dirs[3 * (size_t)grow + 1]; z = dirs[3 * (size_t)grow + 2]; }
        const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
        uint4 s0, s1;
        s0.x = pack2(0.28209479177387814f, -0.48860251190291987f * y);
        s0.y = pack2(0.48860251190291987f * z, -0.48860251190291987f * x);
        s0.z = pack2(1.0925484305920792f * xy, -1.0925484305920792f * yz);
        s0.w = pack2(0.94617469575755997f * z2 - 0.31539156525251999f, -1.0925484305920792f * xz);
        s1.x = pack2(0.54627421529603959f * x2 - 0.54627421529603959f * y2, 0.59004358992664352f * y * (-3.0f * x2 + y2));
        s1.y = pack2(2.8906114426405538f * xy * z, 0.45704579946446572f * y * (1.0f - 5.0f * z2));
        s1.z = pack2(0.3731763325901154f * z * (5.0f * z2 - 3.0f), 0.45704579946446572f * x * (1.0f - 5.0f * z2));
        s1.w = pack2(1.4453057213202769f * z * (x2 - y2), 0.59004358992664352f * x * (-x2 + 3.0f * y2));
        if (!ok) s0 = s1 = make_uint4(0u, 0u, 0u, 0u);
        *reinterpret_cast<uint4 *>(t + ((0u ^ sw) << 4)) = s0;
        *reinterpret_cast<uint4 *>(t + ((1u ^ sw) << 4)) = s1;
    }
    if (parts & 2u) {
        uint4 a = make_uint4(0u, 0u, 0u, 0u), b = a;
        if (ok) { a = __ldg(reinterpret_cast<const uint4 *>(h) + 2 * (size_t)grow); b = __ldg(reinterpret_cast<const uint4 *>(h) + 2 * (size_t)grow + 1); }
        h01 = a.x;
        uint4 g0, g1;  // geo_feat = h[1..15] shifted down by one half, then a zero pad
        g0.x = __byte_perm(a.x, a.y, 0x5432); g0.y = __byte_perm(a.y, a.z, 0x5432); g0.z = __byte_perm(a.z, a.w, 0x5432); g0.w = __byte_perm(a.w, b.x, 0x5432);
        g1.x = __byte_perm(b.x, b.y, 0x5432); g1.y = __byte_perm(b.y, b.z, 0x5432); g1.z = __byte_perm(b.z, b.w, 0x5432); g1.w = b.w >> 16;
        *reinterpret_cast<uint4 *>(t + ((2u ^ sw) << 4)) = g0;
        *reinterpret_cast<uint4 *>(t + ((3u ^ sw) << 4)) = g1;
    }
    return h01;
}
// upstream tile of the colour net's backward: columns [col8*8, col8*8 + 16) of row `row` = [g * rgb (1 - rgb) x 3 | 0 x 13]
__device__ __forceinline__ void nerf_grad_c_row(uint8_t *tile, const nerf_io_t &io, uint32_t grow, bool ok, uint32_t row, uint32_t col8) {
    float g[3] = {0.f, 0.f, 0.f};
    if (ok) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const float r = io.rgbs[3 * (size_t)grow + c];
            g[c] = io.grad_rgbs[3 * (size_t)grow + c] * r * (1.0f - r);
        }
    }
    uint8_t *t = tile + row * kTileRowBytes;
    *reinterpret_cast<uint4 *>(t + (((col8) ^ (row & 7u)) << 4)) = make_uint4(pack2(g[0], g[1]), pack2(g[2], 0.f), 0u, 0u);
    *reinterpret_cast<uint4 *>(t + (((col8 + 1u) ^ (row & 7u)) << 4)) = make_uint4(0u, 0u, 0u, 0u);
}

// ==========================================================================================================
// forward / inference
// ==========================================================================================================
template <int HID> struct split_of { static constexpr int NS = HID >= 64 ? 4 : (HID == 32 ? 2 : 1); };

__device__ __forceinline__ void pipe_sync(uint32_t id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }

// One CTA per SM runs NP independent tile pipelines (one warpgroup of 128 threads each) that share the staged
// weights: measured on B200 a layer step is latency bound (each tcgen05.mma of this size takes ~80-100 cycles to
// ISSUE, the accumulator round trip another ~600), so throughput comes from tiles in flight, not from a
// faster single chain.  Pipeline p: rows = its tile, TMEM columns [p*accw, (p+1)*accw), mbarrier p, named barrier 1+p.
template <typename WT, int HID, bool GACT, int IO>
__global__ void __launch_bounds__(1024) k_ffmlp_fwd(const __half *__restrict__ x, const WT *__restrict__ w, __half *__restrict__ y,
                                                    const mlp_t m, uint32_t ntiles, uint32_t tmem_cols, uint32_t accw,
                                                    const int *__restrict__ limit, const nerf_io_t io) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar[8];
    const uint32_t Beff = eff_batch(m.B, limit);  // device-side batch (ngp_call_opts.batch_limit)
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t np = blockDim.x >> 7, pipe = tid >> 7, row = tid & 127u;
    const uint32_t xbytes = ktiles(m.in) * kActTileBytes, abytes = ktiles(m.hid) * kActTileBytes;
    uint8_t *sw = smem;
    uint8_t *sx = smem + m.wbytes + pipe * (xbytes + abytes);
    uint8_t *sa = sx + xbytes;

    stage_weights<WT>(sw, w, m, tid, blockDim.x);
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid < np) mbar_init(&s_bar[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem + pipe * accw;
    const uint32_t trow = tmem + (((warp & 3u) * 32u) << 16);  // this warp's 32 TMEM lanes
    uint64_t *bar = &s_bar[pipe];
    const uint32_t bid = 1u + pipe;
    uint32_t phase = 0;

    for (uint32_t tile = blockIdx.x * np + pipe; tile < ntiles; tile += gridDim.x * np) {
        if constexpr (IO == 2) {  // x = h [M,16]: thread `row` builds its row of [SH4(dir) | h1..h15 | 0]
            const uint32_t gr = tile * kTileRows + row;
            nerf_color_row(sx, io.dirs, x, gr, gr < Beff, row, 3u);
        } else {
            load_rows_async(sx, x, tile * kTileRows, Beff, m.in, row, 128);
            cp_async_commit();
            cp_async_wait<0>();
        }
        fence_proxy_async();
        pipe_sync(bid);
        if (row == 0) {
            fence_after_sync();
            issue_fwd(tmem, sx, sw + m.soff[0], m.hid, m.in);
            mma_commit(bar);
        }
        for (uint32_t l = 0; l < m.nl; ++l) {
            mbar_wait(bar, phase);
            phase ^= 1u;
            fence_after_sync();
            constexpr int CW = HID < 32 ? HID : 32;  // 1024 threads per CTA: <= 64 registers, drain 32 columns at a time
#pragma unroll
            for (int c0 = 0; c0 < HID; c0 += CW) epi_relu_to_tile<CW>(trow, sa, row, c0, GACT ? m.act : 0u);
            fence_proxy_async();
            fence_before_sync();
            pipe_sync(bid);
            if (row == 0) {
                fence_after_sync();
                issue_fwd(tmem, sa, sw + m.soff[l + 1], m.rows[l + 1], m.hid);
                mma_commit(bar);
            }
        }
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        const uint32_t grow = tile * kTileRows + row;
        if constexpr (IO == 0) {
            epi_to_global(trow, y + (size_t)grow * m.out, 0, m.out, grow < Beff);
        } else {  // out == 16 (host-checked): one 16-column drain
            float v[16];
            tmem_ld16(trow, v);
            if (grow < Beff) {
                if constexpr (IO == 1) {
                    uint4 lo, hi;
                    lo.x = pack2(v[0], v[1]); lo.y = pack2(v[2], v[3]); lo.z = pack2(v[4], v[5]); lo.w = pack2(v[6], v[7]);
                    hi.x = pack2(v[8], v[9]); hi.y = pack2(v[10], v[11]); hi.z = pack2(v[12], v[13]); hi.w = pack2(v[14], v[15]);
                    reinterpret_cast<uint4 *>(y + (size_t)grow * 16u)[0] = lo;
                    reinterpret_cast<uint4 *>(y + (size_t)grow * 16u)[1] = hi;
                    io.sigmas[grow] = io.density_scale * __expf(half_round(v[0]));  // trunc_exp forward of the stored half
                } else {
#pragma unroll
                    for (int c = 0; c < 3; ++c) io.rgbs[3 * (size_t)grow + c] = half_round(1.0f / (1.0f + __expf(-half_round(v[c]))));
                }
            }
        }
        fence_before_sync();
        pipe_sync(bid);  // accumulator drained before this pipeline's next first MMA overwrites it
    }
    __syncthreads();
    if (warp == 0) tmem_dealloc(s_tmem, tmem_cols);
}

// ==========================================================================================================
// both nets of NeRFNetwork(--ff) in ONE launch: enc -> sigma net -> (h, sigma) -> [SH4(dir) | h1..h15 | 0] -> colour net -> rgb
// ==========================================================================================================
// Same pipelines as k_ffmlp_fwd (one warpgroup per 128-row tile, thread = row); the weights of both nets are staged once per
// CTA (44 KB), the X tile that held the encoding becomes the colour net's input tile as soon as the sigma net's first product
// has read it, and the row of h a thread has just drained goes into that tile from registers: h [M,16] is written for the
// backward (TRAIN) but never read back, and in inference it does not exist.  Arithmetic per sample is that of
// ngp_nerf_sigma_forward + ngp_nerf_color_forward, expression for expression.
template <bool TRAIN>
__global__ void __launch_bounds__(1024) k_nerf_nets_fwd(const __half *__restrict__ enc, const float *__restrict__ w_sig, const float *__restrict__ w_col,
                                                        __half *__restrict__ h_out, const mlp_t ms, const mlp_t mc, uint32_t ntiles, uint32_t tmem_cols,
                                                        const int *__restrict__ limit, const nerf_io_t io) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar[8];
    const uint32_t Beff = eff_batch(ms.B, limit);
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t np = blockDim.x >> 7, pipe = tid >> 7, row = tid & 127u;
    uint8_t *sw1 = smem, *sw2 = smem + ms.wbytes;
    uint8_t *sx = smem + ms.wbytes + mc.wbytes + pipe * (2 * kActTileBytes);
    uint8_t *sa = sx + kActTileBytes;
    stage_weights<float>(sw1, w_sig, ms, tid, blockDim.x);
    stage_weights<float>(sw2, w_col, mc, tid, blockDim.x);
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid < np) mbar_init(&s_bar[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem + pipe * 64u;
    const uint32_t trow = tmem + (((warp & 3u) * 32u) << 16);
    uint64_t *bar = &s_bar[pipe];
    const uint32_t bid = 1u + pipe;
    uint32_t phase = 0;
    // one round: wait for the product, drain 64 columns with ReLU into `sa`, hand the next product to the tensor core
    auto hidden_round = [&](const uint8_t *w_next, uint32_t n_next) {
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        epi_relu_to_tile<32>(trow, sa, row, 0u);
        epi_relu_to_tile<32>(trow, sa, row, 32u);
        fence_proxy_async();
        fence_before_sync();
        pipe_sync(bid);
        if (row == 0) {
            fence_after_sync();
            issue_fwd(tmem, sa, w_next, n_next, 64u);
            mma_commit(bar);
        }
    };
    for (uint32_t tile = blockIdx.x * np + pipe; tile < ntiles; tile += gridDim.x * np) {
        const uint32_t grow = tile * kTileRows + row;
        const bool ok = grow < Beff;
        load_rows_async(sx, enc, tile * kTileRows, Beff, 32u, row, 128);
        cp_async_commit();
        float dx = 0.f, dy = 0.f, dz = 0.f;  // the direction of this thread's sample: requested now, used after the sigma net
        if (ok) { dx = io.dirs[3 * (size_t)grow]; dy = io.dirs[3 * (size_t)grow + 1]; dz = io.dirs[3 * (size_t)grow + 2]; }
        cp_async_wait<0>();
        fence_proxy_async();
        pipe_sync(bid);
        if (row == 0) {
            fence_after_sync();
            issue_fwd(tmem, sx, sw1 + ms.soff[0], 64u, 32u);
            mma_commit(bar);
        }
        hidden_round(sw1 + ms.soff[1], 64u);
        hidden_round(sw1 + ms.soff[2], 16u);
        // ---- h, sigma, and the colour net's input tile (into the X tile: the sigma net's layer 0 is long done with it) ----
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        {
            float v[16];
            tmem_ld16(trow, v);
            uint4 lo, hi;
            lo.x = pack2(v[0], v[1]); lo.y = pack2(v[2], v[3]); lo.z = pack2(v[4], v[5]); lo.w = pack2(v[6], v[7]);
            hi.x = pack2(v[8], v[9]); hi.y = pack2(v[10], v[11]); hi.z = pack2(v[12], v[13]); hi.w = pack2(v[14], v[15]);
            if (!ok) lo = hi = make_uint4(0u, 0u, 0u, 0u);
            if (ok) {
                if (TRAIN) {
                    reinterpret_cast<uint4 *>(h_out + (size_t)grow * 16u)[0] = lo;
                    reinterpret_cast<uint4 *>(h_out + (size_t)grow * 16u)[1] = hi;
                }
                io.sigmas[grow] = io.density_scale * __expf(half_round(v[0]));
            }
            const float x = dx, y = dy, z = dz;
            const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
            uint4 s0, s1;
            s0.x = pack2(0.28209479177387814f, -0.48860251190291987f * y);
            s0.y = pack2(0.48860251190291987f * z, -0.48860251190291987f * x);
            s0.z = pack2(1.0925484305920792f * xy, -1.0925484305920792f * yz);
            s0.w = pack2(0.94617469575755997f * z2 - 0.31539156525251999f, -1.0925484305920792f * xz);
            s1.x = pack2(0.54627421529603959f * x2 - 0.54627421529603959f * y2, 0.59004358992664352f * y * (-3.0f * x2 + y2));
            s1.y = pack2(2.8906114426405538f * xy * z, 0.45704579946446572f * y * (1.0f - 5.0f * z2));
            s1.z = pack2(0.3731763325901154f * z * (5.0f * z2 - 3.0f), 0.45704579946446572f * x * (1.0f - 5.0f * z2));
            s1.w = pack2(1.4453057213202769f * z * (x2 - y2), 0.59004358992664352f * x * (-x2 + 3.0f * y2));
            if (!ok) s0 = s1 = make_uint4(0u, 0u, 0u, 0u);
            uint4 g0, g1;  // geo_feat = h[1..15] shifted down by one half, then a zero pad
            g0.x = __byte_perm(lo.x, lo.y, 0x5432); g0.y = __byte_perm(lo.y, lo.z, 0x5432); g0.z = __byte_perm(lo.z, lo.w, 0x5432); g0.w = __byte_perm(lo.w, hi.x, 0x5432);
            g1.x = __byte_perm(hi.x, hi.y, 0x5432); g1.y = __byte_perm(hi.y, hi.z, 0x5432); g1.z = __byte_perm(hi.z, hi.w, 0x5432); g1.w = hi.w >> 16;
            uint8_t *t = sx + row * kTileRowBytes;
            const uint32_t swz = row & 7u;
            *reinterpret_cast<uint4 *>(t + ((0u ^ swz) << 4)) = s0;
            *reinterpret_cast<uint4 *>(t + ((1u ^ swz) << 4)) = s1;
            *reinterpret_cast<uint4 *>(t + ((2u ^ swz) << 4)) = g0;
            *reinterpret_cast<uint4 *>(t + ((3u ^ swz) << 4)) = g1;
        }
        fence_proxy_async();
        fence_before_sync();
        pipe_sync(bid);
        if (row == 0) {
            fence_after_sync();
            issue_fwd(tmem, sx, sw2 + mc.soff[0], 64u, 32u);
            mma_commit(bar);
        }
        for (uint32_t l = 0; l < mc.nl; ++l) hidden_round(sw2 + mc.soff[l + 1], mc.rows[l + 1]);
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        {
            float v[16];
            tmem_ld16(trow, v);
            if (ok) {
#pragma unroll
                for (int c = 0; c < 3; ++c) io.rgbs[3 * (size_t)grow + c] = half_round(1.0f / (1.0f + __expf(-half_round(v[c]))));
            }
        }
        fence_before_sync();
        pipe_sync(bid);  // accumulator drained before this pipeline's next first MMA overwrites it
    }
    __syncthreads();
    if (warp == 0) tmem_dealloc(s_tmem, tmem_cols);
}

// ---- the same two nets with every activation in TENSOR MEMORY ------------------------------------------------------------
// The drain of a layer (tcgen05.ld -> ReLU -> fp16) writes the next layer's A operand back into TMEM with tcgen05.st (row = lane,
// two K elements per 32-bit column) and the products read it there (tcgen05.mma, A from TMEM): no activation tile in shared
// memory at all, the inputs go global -> registers -> TMEM.  Why: measured (benchmarks/umma_issue_bench.cu, tmem_ld_bench.cu) a
// shared-memory-operand MMA holds the SM's 128 B/clk shared-memory port for (4 KB + N x 32 B) / 128 cycles and a 128 x 64 drain
// for 16 KB / 128 = 128 cycles; with A in TMEM only the weights (N x 32 B per instruction) cross that port.  Shared memory
// holds the two weight images (44 KB); a pipeline owns 96 TMEM columns (64 accumulator + 32 operand), five pipelines per SM.
// Same values and the same accumulation order as k_nerf_nets_fwd: results are bit-identical.
__device__ __forceinline__ void issue_fwd_ts(uint32_t d_tmem, uint32_t a_tmem, const uint8_t *w_img, uint32_t N, uint32_t K) {
    const uint32_t idesc = make_idesc(128, N, false, false);
    const uint32_t b_base = smem_u32(w_img);
    for (uint32_t k16 = 0; k16 < K / 16; ++k16) {
        const uint64_t bd = make_desc(b_base + (k16 >> 2) * N * kTileRowBytes + (k16 & 3u) * 32u, 0, kAtomBytes);
        mma_f16_ts(d_tmem, a_tmem + k16 * 8u, bd, idesc, k16 > 0 ? 1u : 0u);
    }
}

// The plain operator (and the sigma net, IO 1) the same way, for hidden <= 64: pipeline columns = accumulator (max(hid, out)
// rounded to 32) + operand (max(in, hid) / 2).  Inputs go global -> registers -> TMEM in 32-half pieces.
// D[128 x N] = G[128 x K] (TMEM, K-major) * W, W = weight image [K rows x N cols] read as an MN-major B operand (see issue_bwd_dx)
__device__ __forceinline__ void issue_bwd_dx_ts(uint32_t d_tmem, uint32_t a_tmem, const uint8_t *w_img, uint32_t N, uint32_t K) {
    const uint32_t idesc = make_idesc(128, N, false, true);
    const uint32_t b_base = smem_u32(w_img);
    const uint32_t b_lbo = N > 64 ? K * kTileRowBytes : 0u;
    for (uint32_t k16 = 0; k16 < K / 16; ++k16)
        mma_f16_ts(d_tmem, a_tmem + k16 * 8u, make_desc(b_base + k16 * 2048u, b_lbo, kAtomBytes), idesc, k16 > 0 ? 1u : 0u);
}

template <typename WT, int HID, bool GACT, int IO>
__global__ void __launch_bounds__(1024) k_ffmlp_fwd_tm(const __half *__restrict__ x, const WT *__restrict__ w, __half *__restrict__ y, const mlp_t m,
                                                       uint32_t ntiles, uint32_t tmem_cols, uint32_t accw, uint32_t pcols,
                                                       const int *__restrict__ limit, const nerf_io_t io) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar[8];
    __shared__ float s_loss;   // IO 3
    float loss_acc = 0.f;      // IO 3: this thread's rows
    const uint32_t Beff = eff_batch(m.B, limit);
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t np = blockDim.x >> 7, pipe = tid >> 7, row = tid & 127u;
    if (IO == 3 && tid == 0) s_loss = 0.f;
    uint8_t *sw = smem;
    stage_weights<WT>(sw, w, m, tid, blockDim.x);
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid < np) mbar_init(&s_bar[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t acc = s_tmem + pipe * pcols, aop = acc + accw;
    const uint32_t lane_off = ((warp & 3u) * 32u) << 16;
    const uint32_t trow = acc + lane_off, arow = aop + lane_off;
    uint64_t *bar = &s_bar[pipe];
    const uint32_t bid = 1u + pipe;
    uint32_t phase = 0;
    auto publish_and_issue = [&](const uint8_t *w_img, uint32_t N, uint32_t K) {
        tmem_st_wait();
        fence_before_sync();
        pipe_sync(bid);
        if (row == 0) {
            fence_after_sync();
            issue_fwd_ts(acc, aop, w_img, N, K);
            mma_commit(bar);
        }
    };
    for (uint32_t tile = blockIdx.x * np + pipe; tile < ntiles; tile += gridDim.x * np) {
        const uint32_t grow = tile * kTileRows + row;
        const bool ok = grow < Beff;
        for (uint32_t c0 = 0; c0 < m.in; c0 += 32) {  // 32 halves = 16 columns per store; in % 32 == 16 leaves a half-filled last piece
            uint32_t p[16];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const bool in_row = ok && c0 + 8u * q < m.in;
                const uint4 v = in_row ? __ldg(reinterpret_cast<const uint4 *>(x + (size_t)grow * m.in + c0) + q) : make_uint4(0u, 0u, 0u, 0u);
                p[4 * q] = v.x; p[4 * q + 1] = v.y; p[4 * q + 2] = v.z; p[4 * q + 3] = v.w;
            }
            tmem_st16(arow + (c0 >> 1), p);
        }
        publish_and_issue(sw + m.soff[0], m.hid, m.in);
        for (uint32_t l = 0; l < m.nl; ++l) {
            mbar_wait(bar, phase);
            phase ^= 1u;
            fence_after_sync();
            constexpr int CW = HID < 32 ? HID : 32;
#pragma unroll
            for (int c0 = 0; c0 < HID; c0 += CW) {
                float v[CW];
                tmem_ldN<CW>(trow + c0, v);
                uint32_t p[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) p[j] = 0u;
#pragma unroll
                for (int j = 0; j < CW / 2; ++j) {
                    if (GACT && m.act != 0u) p[j] = pack2(act_fwd(m.act, v[2 * j]), act_fwd(m.act, v[2 * j + 1]));
                    else p[j] = relu2(v[2 * j], v[2 * j + 1]);
                }
                tmem_st16(arow + (c0 >> 1), p);  // HID 16: the upper eight columns written are zeros past this layer's K, never read
            }
            publish_and_issue(sw + m.soff[l + 1], m.rows[l + 1], m.hid);
        }
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        if constexpr (IO == 0) {
            epi_to_global(trow, y + (size_t)grow * m.out, 0, m.out, ok);
        } else if constexpr (IO == 3) {
            // SDF regression (nerf_step.cu k_sdf_loss, same expressions): the prediction is the fp16 value the plain forward would
            // have stored in column 0; what leaves the SM is the loss gradient row, not the output row
            float v[16];
            tmem_ld16(trow, v);
            if (ok) {
                const float pred = half_round(v[0]);
                const float yt = io.target[grow];
                const float wgt = 1.0f / (fabsf(yt) + 1e-2f), d = pred - yt;
                loss_acc += fabsf(d) * wgt;
                const float s = io.loss_scale ? *io.loss_scale : 1.0f;
                const float g = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
                reinterpret_cast<uint4 *>(y + (size_t)grow * 16u)[0] = make_uint4(pack2(s * g * wgt / (float)Beff, 0.f), 0u, 0u, 0u);
                reinterpret_cast<uint4 *>(y + (size_t)grow * 16u)[1] = make_uint4(0u, 0u, 0u, 0u);
            }
        } else {
            float v[16];
            tmem_ld16(trow, v);
            if (ok) {
                uint4 lo, hi;
                lo.x = pack2(v[0], v[1]); lo.y = pack2(v[2], v[3]); lo.z = pack2(v[4], v[5]); lo.w = pack2(v[6], v[7]);
                hi.x = pack2(v[8], v[9]); hi.y = pack2(v[10], v[11]); hi.z = pack2(v[12], v[13]); hi.w = pack2(v[14], v[15]);
                reinterpret_cast<uint4 *>(y + (size_t)grow * 16u)[0] = lo;
                reinterpret_cast<uint4 *>(y + (size_t)grow * 16u)[1] = hi;
                io.sigmas[grow] = io.density_scale * __expf(half_round(v[0]));
            }
        }
    }
    if constexpr (IO == 3) {  // one global atomic per CTA
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
        if ((tid & 31u) == 0u && loss_acc != 0.f) atomicAdd(&s_loss, loss_acc);
    }
    fence_before_sync();
    __syncthreads();
    if (IO == 3 && tid == 0 && s_loss != 0.f) atomicAdd(io.loss_out, s_loss / (float)Beff);
    if (warp == 0) tmem_dealloc(s_tmem, tmem_cols);
}

template <bool TRAIN>
__global__ void __launch_bounds__(640) k_nerf_nets_fwd_tm(const __half *__restrict__ enc, const float *__restrict__ w_sig, const float *__restrict__ w_col,
                                                          __half *__restrict__ h_out, const mlp_t ms, const mlp_t mc, uint32_t ntiles, uint32_t tmem_cols,
                                                          const int *__restrict__ limit, const nerf_io_t io) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar[8];
    const uint32_t Beff = eff_batch(ms.B, limit);
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t np = blockDim.x >> 7, pipe = tid >> 7, row = tid & 127u;
    uint8_t *sw1 = smem, *sw2 = smem + ms.wbytes;
    stage_weights<float>(sw1, w_sig, ms, tid, blockDim.x);
    stage_weights<float>(sw2, w_col, mc, tid, blockDim.x);
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid < np) mbar_init(&s_bar[tid], 1);
    if (tid == 0) mbar_init_fence();
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t acc = s_tmem + pipe * 96u, aop = acc + 64u;          // accumulator / operand columns of this pipeline
    const uint32_t lane_off = ((warp & 3u) * 32u) << 16;                 // this warp's 32 TMEM lanes
    const uint32_t trow = acc + lane_off, arow = aop + lane_off;
    uint64_t *bar = &s_bar[pipe];
    const uint32_t bid = 1u + pipe;
    uint32_t phase = 0;
    // hand the operand this thread has just stored to the tensor core: every row of the tile must be in TMEM first
    auto publish_and_issue = [&](const uint8_t *w_img, uint32_t N, uint32_t K) {
        tmem_st_wait();
        fence_before_sync();
        pipe_sync(bid);
        if (row == 0) {
            fence_after_sync();
            issue_fwd_ts(acc, aop, w_img, N, K);
            mma_commit(bar);
        }
    };
    auto hidden_round = [&](const uint8_t *w_next, uint32_t n_next) {
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
#pragma unroll
        for (uint32_t c0 = 0; c0 < 64; c0 += 32) {
            float v[32];
            tmem_ld32(trow + c0, v);
            uint32_t p[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) p[j] = relu2(v[2 * j], v[2 * j + 1]);
            tmem_st16(arow + (c0 >> 1), p);
        }
        publish_and_issue(w_next, n_next, 64u);
    };
    for (uint32_t tile = blockIdx.x * np + pipe; tile < ntiles; tile += gridDim.x * np) {
        const uint32_t grow = tile * kTileRows + row;
        const bool ok = grow < Beff;
        {
            uint4 x[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) x[q] = ok ? __ldg(reinterpret_cast<const uint4 *>(enc + (size_t)grow * 32u) + q) : make_uint4(0u, 0u, 0u, 0u);
            const uint32_t p[16] = {x[0].x, x[0].y, x[0].z, x[0].w, x[1].x, x[1].y, x[1].z, x[1].w, x[2].x, x[2].y, x[2].z, x[2].w, x[3].x, x[3].y, x[3].z, x[3].w};
            tmem_st16(arow, p);
        }
        float dx = 0.f, dy = 0.f, dz = 0.f;
        if (ok) { dx = io.dirs[3 * (size_t)grow]; dy = io.dirs[3 * (size_t)grow + 1]; dz = io.dirs[3 * (size_t)grow + 2]; }
        publish_and_issue(sw1 + ms.soff[0], 64u, 32u);
        hidden_round(sw1 + ms.soff[1], 64u);
        hidden_round(sw1 + ms.soff[2], 16u);
        // ---- h, sigma, and the colour net's input row ----
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        {
            float v[16];
            tmem_ld16(trow, v);
            uint4 lo, hi;
            lo.x = pack2(v[0], v[1]); lo.y = pack2(v[2], v[3]); lo.z = pack2(v[4], v[5]); lo.w = pack2(v[6], v[7]);
            hi.x = pack2(v[8], v[9]); hi.y = pack2(v[10], v[11]); hi.z = pack2(v[12], v[13]); hi.w = pack2(v[14], v[15]);
            if (ok) {
                if (TRAIN) {
                    reinterpret_cast<uint4 *>(h_out + (size_t)grow * 16u)[0] = lo;
                    reinterpret_cast<uint4 *>(h_out + (size_t)grow * 16u)[1] = hi;
                }
                io.sigmas[grow] = io.density_scale * __expf(half_round(v[0]));
            } else {
                lo = hi = make_uint4(0u, 0u, 0u, 0u);
            }
            const float x = dx, y = dy, z = dz;
            const float xy = x * y, xz = x * z, yz = y * z, x2 = x * x, y2 = y * y, z2 = z * z;
            uint32_t p[16];
            p[0] = pack2(0.28209479177387814f, -0.48860251190291987f * y);
            p[1] = pack2(0.48860251190291987f * z, -0.48860251190291987f * x);
            p[2] = pack2(1.0925484305920792f * xy, -1.0925484305920792f * yz);
            p[3] = pack2(0.94617469575755997f * z2 - 0.31539156525251999f, -1.0925484305920792f * xz);
            p[4] = pack2(0.54627421529603959f * x2 - 0.54627421529603959f * y2, 0.59004358992664352f * y * (-3.0f * x2 + y2));
            p[5] = pack2(2.8906114426405538f * xy * z, 0.45704579946446572f * y * (1.0f - 5.0f * z2));
            p[6] = pack2(0.3731763325901154f * z * (5.0f * z2 - 3.0f), 0.45704579946446572f * x * (1.0f - 5.0f * z2));
            p[7] = pack2(1.4453057213202769f * z * (x2 - y2), 0.59004358992664352f * x * (-x2 + 3.0f * y2));
            if (!ok) {
#pragma unroll
                for (int j = 0; j < 8; ++j) p[j] = 0u;
            }
            // geo_feat = h[1..15] shifted down by one half, then a zero pad
            p[8] = __byte_perm(lo.x, lo.y, 0x5432); p[9] = __byte_perm(lo.y, lo.z, 0x5432); p[10] = __byte_perm(lo.z, lo.w, 0x5432); p[11] = __byte_perm(lo.w, hi.x, 0x5432);
            p[12] = __byte_perm(hi.x, hi.y, 0x5432); p[13] = __byte_perm(hi.y, hi.z, 0x5432); p[14] = __byte_perm(hi.z, hi.w, 0x5432); p[15] = hi.w >> 16;
            tmem_st16(arow, p);
        }
        publish_and_issue(sw2 + mc.soff[0], 64u, 32u);
        for (uint32_t l = 0; l < mc.nl; ++l) hidden_round(sw2 + mc.soff[l + 1], mc.rows[l + 1]);
        mbar_wait(bar, phase);
        phase ^= 1u;
        fence_after_sync();
        {
            float v[16];
            tmem_ld16(trow, v);
            if (ok) {
#pragma unroll
                for (int c = 0; c < 3; ++c) io.rgbs[3 * (size_t)grow + c] = half_round(1.0f / (1.0f + __expf(-half_round(v[c]))));
            }
        }
        // the next tile's first tcgen05.st / product touch columns whose readers (this drain, the last product) are done:
        // the drain above is complete for this thread, and the pipe_sync inside publish_and_issue orders the other rows' drains
        // before the elected thread issues
    }
    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(s_tmem, tmem_cols);
}

// ==========================================================================================================
// backward (recomputes the forward per tile; dW accumulated in TMEM across the CTA's tiles)
// ==========================================================================================================
struct bwd_cols_t {
    uint32_t acc, dw[kMaxLayers + 1], total;
};
__host__ __device__ inline void make_bwd_cols(const mlp_t &m, bool calc_dx, bwd_cols_t &c) {
    uint32_t a = m.hid;
    if (calc_dx && m.in > a) a = m.in;
    a = (a + 31u) & ~31u;
    c.acc = 0;
    uint32_t o = a;
    for (uint32_t i = 0; i <= m.nl; ++i) {
        c.dw[i] = o;
        o += (m.K[i] + 31u) & ~31u;
    }
    c.total = o;
}

template <typename WT, int HID, bool GACT>
__global__ void __launch_bounds__(128 * split_of<HID>::NS) k_ffmlp_bwd(const __half *__restrict__ grad, const __half *__restrict__ x, const WT *__restrict__ w,
                                                   __half *__restrict__ grad_x, float *__restrict__ ws, void *__restrict__ grad_w,
                                                   int grad_w_f32, const mlp_t m, uint32_t ntiles, uint32_t tmem_cols, int calc_dx,
                                                   uint32_t nbuf, const int *__restrict__ limit) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ uint32_t s_last;
    const uint32_t Beff = eff_batch(m.B, limit);  // device-side batch (ngp_call_opts.batch_limit)
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    __shared__ __align__(8) uint64_t s_bar;
    constexpr int NS = split_of<HID>::NS, CW = HID / NS;
    const uint32_t tid = threadIdx.x, warp = tid >> 5;
    const uint32_t row = tid & 127u, cg = tid >> 7;
    const uint32_t xbytes = ktiles(m.in) * kActTileBytes, abytes = ktiles(m.hid) * kActTileBytes, obytes = ktiles(m.out) * kActTileBytes;
    uint8_t *sw = smem;
    uint8_t *p = smem + m.wbytes;
    // nbuf == 2: input / upstream-gradient tiles are prefetched one tile ahead; nbuf == 1 (wide layers that
    // would not fit otherwise): loaded at the top of each tile
    uint8_t *sx[2] = {p, p + (nbuf - 1) * xbytes};
    p += nbuf * xbytes;
    uint8_t *sgo[2] = {p, p + (nbuf - 1) * obytes};
    p += nbuf * obytes;
    uint8_t *sg[2] = {p, p + abytes};
    p += 2 * abytes;
    uint8_t *sh = p;  // nl activation tile sets
    bwd_cols_t tc;
    make_bwd_cols(m, calc_dx != 0, tc);

    uint32_t tile = blockIdx.x;
    if (tile < ntiles && nbuf == 2) {
        load_rows_async(sx[0], x, tile * kTileRows, Beff, m.in, tid, blockDim.x);
        load_rows_async(sgo[0], grad, tile * kTileRows, Beff, m.out, tid, blockDim.x);
    }
    cp_async_commit();
    stage_weights<WT>(sw, w, m, tid, blockDim.x);
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid == 0) {
        mbar_init(&s_bar, 1);
        mbar_init_fence();
    }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t trow = tmem + (((warp & 3u) * 32u) << 16);
    uint32_t phase = 0, buf = 0, first = 1;

    for (; tile < ntiles; tile += gridDim.x, buf ^= (nbuf - 1u)) {
        if (nbuf == 2) {
            const uint32_t next = tile + gridDim.x;
            if (next < ntiles) {
                load_rows_async(sx[buf ^ 1u], x, next * kTileRows, Beff, m.in, tid, blockDim.x);
                load_rows_async(sgo[buf ^ 1u], grad, next * kTileRows, Beff, m.out, tid, blockDim.x);
            }
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            load_rows_async(sx[0], x, tile * kTileRows, Beff, m.in, tid, blockDim.x);
            load_rows_async(sgo[0], grad, tile * kTileRows, Beff, m.out, tid, blockDim.x);
            cp_async_commit();
            cp_async_wait<0>();
        }
        fence_proxy_async();
        __syncthreads();
        // ---- forward recompute: H_l = ReLU(W_l H_{l-1}) kept in shared memory ----
        if (tid == 0) {
            fence_after_sync();
            issue_fwd(tmem + tc.acc, sx[buf], sw + m.soff[0], m.hid, m.in);
            mma_commit(&s_bar);
        }
        for (uint32_t l = 0; l < m.nl; ++l) {
            mbar_wait(&s_bar, phase);
            phase ^= 1u;
            fence_after_sync();
            epi_relu_to_tile<CW>(trow + tc.acc, sh + l * abytes, row, cg * CW, GACT ? m.act : 0u);
            fence_proxy_async();
            fence_before_sync();
            __syncthreads();
            if (tid == 0) {
                fence_after_sync();
                if (l + 1 < m.nl) {
                    issue_fwd(tmem + tc.acc, sh + l * abytes, sw + m.soff[l + 1], m.hid, m.hid);
                } else {  // start of the backward: dH_{nl-1} = dY W_last ; dW_last += dY^T H_{nl-1}
                    issue_bwd_dx(tmem + tc.acc, sgo[buf], sw + m.soff[m.nl], m.hid, m.out);
                }
                mma_commit(&s_bar);
                if (l + 1 == m.nl) issue_dw(tmem + tc.dw[m.nl], sgo[buf], sh + l * abytes, m.out, m.hid, first ? 0u : 1u);
            }
        }
        // ---- backward through the hidden layers ----
        for (int l = (int)m.nl - 1; l >= 0; --l) {
            uint8_t *g = sg[l & 1];
            mbar_wait(&s_bar, phase);
            phase ^= 1u;
            fence_after_sync();
            epi_mask_to_tile<CW>(trow + tc.acc, sh + (uint32_t)l * abytes, g, row, cg * CW, GACT ? m.act : 0u);  // G_l = dH_l * (H_l > 0)
            fence_proxy_async();
            fence_before_sync();
            __syncthreads();
            if (tid == 0) {
                fence_after_sync();
                if (l > 0) {
                    issue_bwd_dx(tmem + tc.acc, g, sw + m.soff[l], m.hid, m.hid);  // dH_{l-1} = G_l W_l
                    mma_commit(&s_bar);
                    issue_dw(tmem + tc.dw[l], g, sh + (uint32_t)(l - 1) * abytes, m.hid, m.hid, first ? 0u : 1u);
                } else {
                    if (calc_dx) issue_bwd_dx(tmem + tc.acc, g, sw + m.soff[0], m.in, m.hid);  // dX = G_0 W_0
                    issue_dw(tmem + tc.dw[0], g, sx[buf], m.hid, m.in, first ? 0u : 1u);
                    mma_commit(&s_bar);
                }
            }
        }
        mbar_wait(&s_bar, phase);
        phase ^= 1u;
        fence_after_sync();
        if (calc_dx) {
            const uint32_t grow = tile * kTileRows + row;
            for (uint32_t c0 = cg * 16u; c0 < m.in; c0 += NS * 16u)
                epi_to_global(trow + tc.acc, grad_x + (size_t)grow * m.in, c0, c0 + 16u, grow < Beff);
        }
        fence_before_sync();
        __syncthreads();
        first = 0;
    }
    cp_async_wait<0>();

    // ---- flush this CTA's dW accumulators: TMEM lane = output row of the matrix ----
    if (!first) {
        for (uint32_t i = 0; i <= m.nl; ++i) {
            const uint32_t rows = m.rows[i], K = m.K[i];
            for (uint32_t c0 = cg * 16u; c0 < K; c0 += NS * 16u) {  // warp-uniform trip count: tcgen05.ld is .sync.aligned
                float v[16];
                tmem_ld16(trow + tc.dw[i] + c0, v);
                if (row < rows) {
                    float *d = ws + m.woff[i] + (size_t)row * K + c0;
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
    // ---- last CTA: fp32 sum -> grad_weights (requested dtype); leave the workspace zeroed for the next call ----
    uint32_t *counter = reinterpret_cast<uint32_t *>(ws + ((m.nweights + 3u) & ~3u));
    if (tid == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (uint32_t i = tid; i < m.nweights; i += blockDim.x) {
        const float v = __ldcg(ws + i);
        if (grad_w_f32) reinterpret_cast<float *>(grad_w)[i] = v;
        else reinterpret_cast<__half *>(grad_w)[i] = __float2half_rn(v);
        ws[i] = 0.f;
    }
    if (tid == 0) *counter = 0u;
}


// ==========================================================================================================
// backward, multi-pipeline variant
// ==========================================================================================================
// The single-tile kernel above serialises, per tile, 2*nl+1 round trips (MMA issue -> commit -> TMEM drain ->
// st.shared -> barrier); measured on B200 it spends ~200 cycles of SM time per tcgen05.mma against a 32-cycle
// floor.  Here one CTA per SM keeps NP tiles in flight, and NP warpgroups of 128*ns threads each own one tile: they
// drain the accumulator, apply ReLU / the ReLU mask and hand the operands of the next layer step back through an
// mbarrier.  Per pipeline: full[p] (operands in shared memory) -> issuer -> done[p] (tcgen05.commit: accumulator
// ready).  G_l overwrites H_l in place (the commits of a round cover the dW MMA that still reads H_l), which is what
// makes room for 3-4 tiles in 227 KB.
// Two issuer warps (benchmarks/umma_issue_bench.cu, handshake_bench.cu: one thread cannot issue tcgen05.mma faster
// than one per ~64-74 cycles whatever N is, while the pipe takes 48 cycles at N=64 and 40 at N<=32; and every
// hand-off costs the issuer ~250 cycles of polling): warp 0 issues the layer products (H_l, dH_l) of all pipelines,
// warp 1 the weight-gradient products.  The dW accumulators in TMEM are shared by all pipelines and touched by
// warp 1's elected lane only -- one thread's tcgen05.mma stream is ordered.  Each issuer has its own full/done
// barrier pair per pipeline; the tile's threads wait for both before they overwrite an operand.
#ifdef NGP_FFMLP_TRACE
// developer build only (benchmarks/ffmlp_trace.py): CTA 0 logs (clock, tag) pairs of its issuers and of the first warp of each pipeline
__device__ unsigned long long g_tr[8][1024];
__device__ unsigned g_trn[8];
#define NGP_TR(slot, tag)                                                                   \
    do {                                                                                    \
        if (blockIdx.x == 0 && (threadIdx.x & 31u) == 0 && tr_n < 1024) {                   \
            g_tr[slot][tr_n] = ((unsigned long long)clock64() << 12) | (unsigned long long)((tag) & 0xfffu); \
            g_trn[slot] = ++tr_n;  /* stores only: a load of the counter would put ~700 cycles into the timeline */ \
        }                                                                                   \
    } while (0)
#else
#define NGP_TR(slot, tag) do { } while (0)
#endif

template <typename WT, int HID, bool GACT, int IO>
__global__ void __launch_bounds__(64 + 128 * 6) k_ffmlp_bwd_mp(const __half *__restrict__ grad, const __half *__restrict__ x, const WT *__restrict__ w,
                                                               __half *__restrict__ grad_x, float *__restrict__ ws, void *__restrict__ grad_w,
                                                               int grad_w_f32, const mlp_t m, uint32_t ntiles, uint32_t tmem_cols, int calc_dx,
                                                               uint32_t np, uint32_t ns, uint32_t accw, int dual, const int *__restrict__ limit,
                                                               const nerf_io_t io, uint32_t opc) {
    // accw: TMEM columns per pipeline = accumulator + opc operand columns.  opc = 32 (option ffmlp_bwd_tmem; hidden 64, ReLU,
    // self-issuing pipelines; OFF by default -- A/B in one job: 218 vs 219 ps/sample for the colour net, the backward is bound by
    // its per-round chain (drain -> barrier -> issue behind the other pipelines' and the weight-gradient products -> completion),
    // not by operand traffic; the forward kernels, which have no weight-gradient products and need no shared-memory copy of the
    // activations at all, gained 36 % from the same idea):
    // the drains also store H_l / G_l into the pipeline's operand columns and the K = 64 layer products read A from there
    // (tcgen05.mma, A in TMEM) -- 4 KB less shared-memory operand traffic per instruction on the port that bounds this kernel;
    // the products whose A is the input tile or the upstream tile (K = in, K = out) and all weight-gradient products (both
    // operands transposed) keep their shared-memory operands.  opc = 0: every operand in shared memory.
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ uint32_t s_last;
    __shared__ __align__(8) uint64_t s_full[4], s_done[4];    // layer products (single-issuer mode: everything)
    __shared__ __align__(8) uint64_t s_full2[4], s_done2[4];  // weight-gradient products (two-issuer mode)
    constexpr int STEP = HID >= 32 ? 32 : 16;  // accumulator columns drained per tcgen05.ld
    const uint32_t cw = m.hid / ns;            // columns of a row owned by one of the ns threads that share it
    const uint32_t Beff = eff_batch(m.B, limit);
    ntiles = min(ntiles, div_round_up(Beff, kTileRows));
    const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const uint32_t xbytes = ktiles(m.in) * kActTileBytes, abytes = ktiles(m.hid) * kActTileBytes, obytes = ktiles(m.out) * kActTileBytes;
    // narrow nets: dY [128 x out] rides in the unused columns of the X tile ([X | dY | pad], in + out <= 64)
    const bool packed = m.in + m.out <= 64u;
    const uint32_t goff = packed ? m.in * 2u : xbytes;  // byte offset of the dY operand from the X tile
    const uint32_t per_pipe = xbytes + (packed ? 0u : obytes) + m.nl * abytes;
    const uint32_t hoff = per_pipe - m.nl * abytes;
    uint8_t *sw = smem;
    uint32_t dwc[kMaxLayers + 1];
    {
        uint32_t o = np * accw;
        for (uint32_t i = 0; i <= m.nl; ++i) {
            dwc[i] = o;
            o += (m.K[i] + 31u) & ~31u;
        }
    }
    stage_weights<WT>(sw, w, m, tid, blockDim.x);
    if (warp == 0) tmem_alloc(&s_tmem, tmem_cols);
    if (tid < np) {
        mbar_init(&s_full[tid], 4 * ns);
        mbar_init(&s_done[tid], 1);
        mbar_init(&s_full2[tid], 4 * ns);
        mbar_init(&s_done2[tid], 1);
    }
    if (tid == 0) mbar_init_fence();
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t stride = gridDim.x * np;
#ifdef NGP_FFMLP_DEBUG
    const bool abl_dw = (dual >> 4) & 1, abl_mm = (dual >> 5) & 1;  // timing ablations (developer builds, option ffmlp_ablate): wrong results
#else
    constexpr bool abl_dw = false, abl_mm = false;
#endif
    // self-issue: every pipeline issues its own layer products (one elected thread of its first warp, after a named
    // barrier over the pipeline's threads) instead of handing them to warp 0.  The central issuer's control code
    // (poll, bookkeeping, ~1.2k cycles per service, three pipelines served in turn) was the critical path of the
    // kernel: with every MMA removed it still ran at 74 % of its full time (benchmarks/ffmlp_trace.py).  Weight
    // gradients stay with warp 1: their TMEM accumulators are shared by the pipelines and one thread's MMA stream is ordered.
    const bool selfi = (dual >> 1) & 1;
    dual &= 1;
#ifdef NGP_FFMLP_TRACE
    unsigned tr_n = 0;
    bool tr_hit = true;
#endif

    // ---- an MMA issuer.  ROLE 0: every product; 1: layer products only; 2: weight-gradient products only ----
    auto issuer = [&](auto role_c) {
        constexpr int ROLE = decltype(role_c)::value;
        uint64_t *fullb = ROLE == 2 ? s_full2 : s_full, *doneb = ROLE == 2 ? s_done2 : s_done;
        const uint32_t r0 = ROLE == 2 ? m.nl : 0u, r1 = 2 * m.nl + 1;  // rounds of a tile this issuer takes part in
        uint32_t left[4], rnd[4], phbits = 0, alive = 0, started = 0;
        for (uint32_t p = 0; p < 4; ++p) {
            const uint32_t t0 = blockIdx.x * np + p;
            left[p] = (p < np && t0 < ntiles) ? (ntiles - t0 + stride - 1) / stride : 0u;
            rnd[p] = r0;
            if (left[p]) alive |= 1u << p;
        }
        while (alive) {
            // one poll for all pipelines: lane p tests full[p].  The ballot keeps everything below warp-uniform on
            // purpose: the descriptors must stay in uniform registers, or every tcgen05.mma pays a register ->
            // uniform-register broadcast loop (measured: ~250 instead of ~64 cycles/MMA)
            const uint32_t pl = lane & 3u;
#ifdef NGP_FFMLP_TRACE
            if (tr_hit) NGP_TR(ROLE == 2 ? 1 : 0, 0xA00u);
            tr_hit = false;
#endif
            const bool ready = lane < 4u && ((alive >> pl) & 1u) && mbar_test(&fullb[pl], (phbits >> pl) & 1u);
            uint32_t rdy = __ballot_sync(0xffffffffu, ready);
            if (rdy == 0u) {
                // nothing to issue: leave the issue slots of this SM sub-partition to the epilogue warps it hosts (the bare
                // spin was ~15 % of all warp instructions of the kernel); the weight-gradient products are one round off
                // every tile's critical path
                __nanosleep(40);
                continue;
            }
            phbits ^= rdy;
#ifdef NGP_FFMLP_TRACE
            tr_hit = true;
            NGP_TR(ROLE == 2 ? 1 : 0, 0xB00u | rdy);
#endif
            fence_after_sync();
            NGP_TR(ROLE == 2 ? 1 : 0, 0x100u | rdy);
            // p stays a run-time (uniform) value: one copy of the issue code instead of four keeps the issuer's
            // working set in the instruction cache (the unrolled version was ~200 KB of SASS and paid ~500 cycles
            // of instruction fetch per round)
            while (rdy) {
                const uint32_t p = (uint32_t)__ffs((int)rdy) - 1u;
                rdy &= rdy - 1u;
                const uint32_t rr = p == 0 ? rnd[0] : p == 1 ? rnd[1] : p == 2 ? rnd[2] : rnd[3];
                uint8_t *sx = smem + m.wbytes + p * per_pipe, *sgo = sx + goff, *sh = sx + hoff;
                const uint32_t acc = tmem + p * accw;
                const bool leader = elect_one();
                if (rr == 0) {
                    if (ROLE != 2 && leader && !abl_mm) issue_fwd(acc, sx, sw + m.soff[0], m.hid, m.in);
                } else if (rr < m.nl) {
                    if (ROLE != 2 && leader && !abl_mm) issue_fwd(acc, sh + (rr - 1) * abytes, sw + m.soff[rr], m.hid, m.hid);
                } else if (rr == m.nl) {  // dH_{nl-1} = dY W_last ; dW_last += dY^T H_{nl-1}
                    if (leader) {
                        if (ROLE != 2 && !abl_mm) issue_bwd_dx(acc, sgo, sw + m.soff[m.nl], m.hid, m.out);
                        if (ROLE != 1 && !abl_dw) issue_dw(tmem + dwc[m.nl], sgo, sh + (m.nl - 1) * abytes, m.out, m.hid, (started >> m.nl) & 1u);
                    }
                    started |= 1u << m.nl;
                } else {
                    const uint32_t l = 2 * m.nl - rr;  // G_l sits where H_l was
                    uint8_t *g = sh + l * abytes;
                    if (leader) {
                        if (l > 0) {
                            if (ROLE != 2 && !abl_mm) issue_bwd_dx(acc, g, sw + m.soff[l], m.hid, m.hid);
                            if (ROLE != 1 && !abl_dw) issue_dw(tmem + dwc[l], g, sh + (l - 1) * abytes, m.hid, m.hid, (started >> l) & 1u);
                        } else {
                            if (ROLE != 2 && calc_dx && !abl_mm) issue_bwd_dx(acc, g, sw + m.soff[0], m.in, m.hid);
                            if (ROLE != 1 && !abl_dw) issue_dw(tmem + dwc[0], g, sx, m.hid, m.in, started & 1u);
                        }
                    }
                    started |= 1u << l;
                }
                NGP_TR(ROLE == 2 ? 1 : 0, 0xC00u | (p << 4) | rr);
                if (leader) mma_commit(&doneb[p]);
                NGP_TR(ROLE == 2 ? 1 : 0, 0x200u | (p << 4) | rr);
                const bool wrap = rr + 1u == r1;
                bool last = false;
#pragma unroll
                for (uint32_t q = 0; q < 4; ++q) {
                    if (q == p) {
                        rnd[q] = wrap ? r0 : rr + 1u;
                        if (wrap) last = --left[q] == 0u;
                    }
                }
                if (last) alive &= ~(1u << p);
            }
            __syncwarp();
        }
    };

    if (warp == 0) {
        if (selfi) { /* layer products are issued by the pipelines */ }
        else if (dual) issuer(std::integral_constant<int, 1>{});
        else issuer(std::integral_constant<int, 0>{});
    } else if (warp == 1) {
        if (dual) issuer(std::integral_constant<int, 2>{});
    } else {
        // ---- tile pipelines ----
        // pipeline p = 4*ns consecutive warps; warp w drains TMEM lanes 32*(w%4).. (hardware rule) = tile rows, and the
        // column group cg = (warp within pipeline) / 4 of width cw, so one row is handled by ns threads in parallel
        const uint32_t wpp = 4u * ns, p = (warp - 2) / wpp, wp = (warp - 2) - p * wpp, cg = wp >> 2, q = warp & 3u;
        const uint32_t row = q * 32u + lane, ptid = wp * 32u + lane, nth = 128u * ns;
        if (p < np) {
            uint8_t *sx = smem + m.wbytes + p * per_pipe, *sh = sx + hoff;
            const uint32_t trow = tmem + p * accw + ((q * 32u) << 16);
            const uint32_t arow = trow + (accw - opc);  // this row's operand columns (opc != 0)
            const bool atm = opc != 0u;
            uint32_t phase = 0, phase2 = 0;
            // round rr of this pipeline's tile: operands are in shared memory (every writer has fenced), accumulator drained
            auto self_issue = [&](uint32_t rr) {
                asm volatile("bar.sync %0, %1;" ::"r"(1u + p), "r"(nth) : "memory");
                if (wp != 0) return;
                fence_after_sync();
                // REDUX puts the pipeline index into a uniform register: everything derived from it (descriptors, TMEM
                // address) stays on the uniform datapath the tcgen05 instructions read from
                const uint32_t pu = __reduce_max_sync(0xffffffffu, p), ru = __reduce_max_sync(0xffffffffu, rr);
                uint8_t *ux = smem + m.wbytes + pu * per_pipe, *ugo = ux + goff, *uh = ux + hoff;
                const uint32_t acc = tmem + pu * accw, aop = acc + (accw - opc);
                if (elect_one()) {
                    if (!abl_mm) {
                        if (ru == 0) issue_fwd(acc, ux, sw + m.soff[0], m.hid, m.in);
                        else if (ru < m.nl) {
                            if (opc) issue_fwd_ts(acc, aop, sw + m.soff[ru], m.hid, m.hid);
                            else issue_fwd(acc, uh + (ru - 1) * abytes, sw + m.soff[ru], m.hid, m.hid);
                        } else if (ru == m.nl) issue_bwd_dx(acc, ugo, sw + m.soff[m.nl], m.hid, m.out);
                        else {
                            const uint32_t l = 2 * m.nl - ru;
                            if (l > 0) {
                                if (opc) issue_bwd_dx_ts(acc, aop, sw + m.soff[l], m.hid, m.hid);
                                else issue_bwd_dx(acc, uh + l * abytes, sw + m.soff[l], m.hid, m.hid);
                            } else if (calc_dx) {
                                if (opc) issue_bwd_dx_ts(acc, aop, sw + m.soff[0], m.in, m.hid);
                                else issue_bwd_dx(acc, uh, sw + m.soff[0], m.in, m.hid);
                            }
                        }
                    }
                    mma_commit(&s_done[pu]);
                }
                __syncwarp();
            };
            for (uint32_t tile = blockIdx.x * np + p; tile < ntiles; tile += stride) {
                uint32_t h01 = 0u;  // IO 2: (h0, h1) of this thread's row, kept for the trunc_exp gradient in the last drain
                if constexpr (IO == 2) {
                    // x = h [M,16], packed layout (in 32 + out 16 <= 64): thread (row, cg) builds its share of row `row`:
                    // [SH4(dir) | h1..h15 | 0 | grad_c 16] -- one thread per row does all of it, two split SH + upstream / geo
                    const uint32_t gr = tile * kTileRows + row;
                    const bool ok = gr < Beff;
                    const uint32_t parts = ns == 1u ? 3u : (cg == 0u ? 1u : 2u);
                    h01 = nerf_color_row(sx, io.dirs, x, gr, ok, row, parts);
                    if (parts & 1u) nerf_grad_c_row(sx, io, gr, ok, row, m.in >> 3);
                } else {
                    load_rows_async(sx, x, tile * kTileRows, Beff, m.in, ptid, nth);
                    if (packed) load_rows_async(sx, grad, tile * kTileRows, Beff, m.out, ptid, nth, m.in >> 3);
                    else load_rows_async(sx + xbytes, grad, tile * kTileRows, Beff, m.out, ptid, nth);
                    if (wp == 0) NGP_TR(2 + p, 0x300u);
                    cp_async_commit();
                    cp_async_wait<0>();
                }
                fence_proxy_async();
                __syncwarp();
                if (selfi) self_issue(0u);
                else if (lane == 0) mbar_arrive(&s_full[p]);  // one arrival per warp
                if (wp == 0) NGP_TR(2 + p, 0x301u);
                for (uint32_t l = 0; l < m.nl; ++l) {    // forward recompute: H_l = ReLU(W_l H_{l-1})
                    mbar_wait(&s_done[p], phase);
                    phase ^= 1u;
                    fence_after_sync();
                    if (wp == 0) NGP_TR(2 + p, 0x400u | l);
                    for (uint32_t c0 = cg * cw; c0 < (cg + 1u) * cw; c0 += STEP)
                        epi_relu_to_tile<STEP>(trow, sh + l * abytes, row, c0, GACT ? m.act : 0u, atm && l + 1 < m.nl, arow);
                    if (atm) tmem_st_wait();
                    fence_proxy_async();
                    fence_before_sync();
                    __syncwarp();
                    if (lane == 0) {
                        if (!selfi) mbar_arrive(&s_full[p]);
                        if (dual && l + 1 == m.nl) mbar_arrive(&s_full2[p]);  // dW_last reads dY and H_{nl-1}
                    }
                    if (selfi) self_issue(l + 1u);
                    if (wp == 0) NGP_TR(2 + p, 0x500u | l);
                }
                for (int l = (int)m.nl - 1; l >= 0; --l) {  // G_l = dH_l * (H_l > 0), in place
                    mbar_wait(&s_done[p], phase);
                    phase ^= 1u;
                    if (dual) {  // the dW product of the previous round still read H_l
                        mbar_wait(&s_done2[p], phase2);
                        phase2 ^= 1u;
                    }
                    fence_after_sync();
                    if (wp == 0) NGP_TR(2 + p, 0x600u | (uint32_t)l);
                    uint8_t *h = sh + (uint32_t)l * abytes;
                    for (uint32_t c0 = cg * cw; c0 < (cg + 1u) * cw; c0 += STEP)
                        epi_mask_to_tile<STEP>(trow, h, h, row, c0, GACT ? m.act : 0u, atm && (l > 0 || calc_dx), arow);
                    if (atm) tmem_st_wait();
                    fence_proxy_async();
                    fence_before_sync();
                    __syncwarp();
                    if (lane == 0) {
                        if (!selfi) mbar_arrive(&s_full[p]);
                        if (dual) mbar_arrive(&s_full2[p]);
                    }
                    if (selfi) self_issue(2u * m.nl - (uint32_t)l);
                    if (wp == 0) NGP_TR(2 + p, 0x700u | (uint32_t)l);
                }
                mbar_wait(&s_done[p], phase);
                phase ^= 1u;
                fence_after_sync();
                if (wp == 0) NGP_TR(2 + p, 0x800u);
                if constexpr (IO == 2) {
                    // grad_h [M,16] = [grad_sigma * density_scale * exp(clamp(h0, -15, 15)) | dL/dcin columns 16..30]; the SH columns'
                    // gradient (dL/ddir) is not needed.  The thread that built the geo part of the row holds h0.
                    if ((ns == 1u || cg == 1u)) {  // warp-uniform: cg is a per-warp value
                        float v[16];
                        tmem_ld16(trow + 16u, v);
                        const uint32_t grow = tile * kTileRows + row;
                        if (grow < Beff) {
                            const float h0 = __half22float2(*reinterpret_cast<const __half2 *>(&h01)).x;
                            const float g0 = io.grad_sigmas[grow] * io.density_scale * __expf(fminf(fmaxf(h0, -15.f), 15.f));
                            uint4 o0, o1;
                            o0.x = pack2(g0, v[0]); o0.y = pack2(v[1], v[2]); o0.z = pack2(v[3], v[4]); o0.w = pack2(v[5], v[6]);
                            o1.x = pack2(v[7], v[8]); o1.y = pack2(v[9], v[10]); o1.z = pack2(v[11], v[12]); o1.w = pack2(v[13], v[14]);
                            reinterpret_cast<uint4 *>(grad_x + (size_t)grow * 16u)[0] = o0;
                            reinterpret_cast<uint4 *>(grad_x + (size_t)grow * 16u)[1] = o1;
                        }
                    }
                } else if (calc_dx) {
                    const uint32_t grow = tile * kTileRows + row;
                    for (uint32_t c0 = cg * 16u; c0 < m.in; c0 += ns * 16u)
                        epi_to_global(trow, grad_x + (size_t)grow * m.in, c0, c0 + 16u, grow < Beff);
                }
                if (dual) {  // dW_0 reads G_0 and X: both are overwritten by the next tile
                    mbar_wait(&s_done2[p], phase2);
                    phase2 ^= 1u;
                }
                if (wp == 0) NGP_TR(2 + p, 0x900u);
                fence_before_sync();
            }
        }
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();

    // ---- flush the CTA's dW accumulators: TMEM lane = output row of the matrix; pipeline index = column group ----
    if (warp > 1 && blockIdx.x * np < ntiles) {
        const uint32_t grp = (warp - 2) >> 2, ngrp = np * ns, q = warp & 3u, row = q * 32u + lane;
        if (grp < ngrp) {
            const uint32_t trow = tmem + ((q * 32u) << 16);
            for (uint32_t i = 0; i <= m.nl; ++i) {
                const uint32_t rows = m.rows[i], K = m.K[i];
                for (uint32_t c0 = grp * 16u; c0 < K; c0 += ngrp * 16u) {  // warp-uniform trip count: tcgen05.ld is .sync.aligned
                    float v[16];
                    tmem_ld16(trow + dwc[i] + c0, v);
                    if (row < rows) {
                        float *d = ws + m.woff[i] + (size_t)row * K + c0;
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(d + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]), "f"(v[j + 3]) : "memory");
                    }
                }
            }
        }
    }
    fence_before_sync();
    __threadfence();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, tmem_cols);
    // ---- last CTA: fp32 sum -> grad_weights (requested dtype); leave the workspace zeroed for the next call ----
    uint32_t *counter = reinterpret_cast<uint32_t *>(ws + ((m.nweights + 3u) & ~3u));
    if (tid == 0) s_last = (atomicAdd(counter, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (uint32_t i = tid; i < m.nweights; i += blockDim.x) {
        const float v = __ldcg(ws + i);
        if (grad_w_f32) reinterpret_cast<float *>(grad_w)[i] = v;
        else reinterpret_cast<__half *>(grad_w)[i] = __float2half_rn(v);
        ws[i] = 0.f;
    }
    if (tid == 0) *counter = 0u;
}

}  // namespace
}  // namespace ngp

// ==========================================================================================================
// host side
// ==========================================================================================================
namespace ngp {
namespace {

constexpr size_t kMaxSmem = 227 * 1024;

int check_dims(const char *fn, uint32_t in, uint32_t out, uint32_t hid, uint32_t nl, uint32_t act, uint32_t out_act, int wdtype) {
    if (!(hid == 16 || hid == 32 || hid == 64 || hid == 128 || hid == 256)) {
        set_error("hidden_dim should in [16, 32, 64, 128, 256]");
        return NGP_EINVAL;
    }
    NGP_CHECK_ARG(in > 0 && in % 16 == 0 && in <= (hid == 256 ? 256u : 128u), "%s: input_dim should be 16 * m, 0 < m <= 8, but got %u", fn, in);
    NGP_CHECK_ARG(out > 0 && out % 16 == 0 && out <= 128, "%s: (padded) output_dim should be 16 * m, 0 < m <= 8, but got %u", fn, out);
    NGP_CHECK_ARG(nl >= 2 && nl <= kMaxLayers, "%s: num_layers should be in [2, %u], but got %u", fn, kMaxLayers, nl);
    NGP_CHECK_ARG(wdtype == NGP_DTYPE_F16 || wdtype == NGP_DTYPE_F32, "%s: weights must be a half or float tensor", fn);
    if (act > 6 || out_act != 6) {
        set_error("%s: activation must be one of the reference's codes 0..6 and output_activation None (6) (got %u, %u)", fn, act, out_act);
        return NGP_EUNSUPPORTED;
    }
    return NGP_OK;
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
    NGP_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return NGP_OK;
}

// io_mode 0: the plain operator.  1 / 2: the sigma / colour net of network_ff with its glue fused in (nerf_io_t; hidden 64, ReLU,
// fp32 parameters, out 16 -- what NeRFNetwork(--ff) uses; anything else is NGP_EUNSUPPORTED)
int check_io(const char *fn, int io_mode, uint32_t in, uint32_t out, uint32_t hid, uint32_t act, int wdtype) {
    if (io_mode == 0) return NGP_OK;
    if (!(hid == 64 && in == 32 && out == 16 && act == 0 && wdtype == NGP_DTYPE_F32)) {
        set_error("%s: the fused network_ff nets are 32 -> 64 x n -> 16, ReLU, fp32 parameters (got in %u hidden %u out %u act %u)", fn, in, hid, out, act);
        return NGP_EUNSUPPORTED;
    }
    return NGP_OK;
}

int run_forward(const char *fn, const void *inputs, const void *weights, int wdtype, uint32_t B, uint32_t in, uint32_t out, uint32_t hid,
                uint32_t nl, uint32_t act, uint32_t out_act, void *forward_buffer, void *outputs, void *stream, int io_mode = 0,
                const nerf_io_t *iop = nullptr) {
    if (B == 0) return NGP_OK;
    NGP_CHECK_ARG(inputs && weights && (outputs || io_mode == 2), "%s: null pointer argument", fn);
    int rc = check_dims(fn, in, out, hid, nl, act, out_act, wdtype);
    if (rc != NGP_OK) return rc;
    rc = check_io(fn, io_mode, in, out, hid, act, wdtype);
    if (rc != NGP_OK) return rc;
    const nerf_io_t io = iop ? *iop : nerf_io_t{};
    if (hid == 256) return wide_forward(fn, inputs, weights, wdtype, B, in, out, nl, act, forward_buffer, outputs, (cudaStream_t)stream);
    mlp_t m;
    make_mlp(m, B, in, out, hid, nl);
    m.act = act;
    const size_t per_pipe = ((size_t)ktiles(in) + ktiles(hid)) * kActTileBytes;
    uint32_t accw = hid > out ? hid : out;
    accw = (accw + 31u) & ~31u;
    const uint32_t ntiles = div_round_up(B, kTileRows);
    if (m.wbytes + per_pipe + 1024 > kMaxSmem) { set_error("%s: configuration needs %zu bytes of shared memory (> %zu)", fn, m.wbytes + per_pipe + 1024, kMaxSmem); return NGP_EUNSUPPORTED; }
    uint32_t np = (uint32_t)((kMaxSmem - 1024 - m.wbytes) / per_pipe);
    if (np > 512 / accw) np = 512 / accw;
    if (np > 8) np = 8;
    const int np_opt = get_option("ffmlp_fwd_pipes", 0);
    if (np_opt > 0 && (uint32_t)np_opt < np) np = (uint32_t)np_opt;
    const uint32_t per_cta_need = div_round_up(ntiles, (uint32_t)sm_count());  // tiles each SM gets anyway
    if (np > per_cta_need) np = per_cta_need;
    const size_t smem = m.wbytes + np * per_pipe;
    const uint32_t cols = tmem_cols_pow2(np * accw);
    uint32_t grid = div_round_up(ntiles, np);
    if (grid > (uint32_t)sm_count()) grid = (uint32_t)sm_count();
    cudaStream_t st = (cudaStream_t)stream;
    // ---- activations in tensor memory (hidden <= 64, plain operator and the sigma net): the default ----
    if (hid <= 64 && io_mode != 2 && get_option("ffmlp_fwd_tmem", 1)) {
        const uint32_t acols = ((in > hid ? in : hid) / 2u + 7u) & ~7u;
        uint32_t opcols = (acols + 15u) & ~15u;              // tcgen05.st writes 16 columns at a time
        if (hid == 16 && opcols < 16u) opcols = 16u;
        const uint32_t pcols = accw + opcols;
        uint32_t npt = 512u / pcols;
        if (npt > 8) npt = 8;
        if (np_opt > 0 && (uint32_t)np_opt < npt) npt = (uint32_t)np_opt;
        if (npt > per_cta_need) npt = per_cta_need;
        if (npt >= 1 && m.wbytes + 1024 <= kMaxSmem) {
            const size_t smem_t = m.wbytes;
            const uint32_t cols_t = tmem_cols_pow2(npt * pcols);
            uint32_t grid_t = div_round_up(ntiles, npt);
            if (grid_t > (uint32_t)sm_count()) grid_t = (uint32_t)sm_count();
#define NGP_FWDTM_LAUNCH(WT_, HID_, G_, IO_)                                                                                 \
    {                                                                                                                        \
        rc = set_smem(k_ffmlp_fwd_tm<WT_, HID_, G_, IO_>, smem_t);                                                           \
        if (rc != NGP_OK) return rc;                                                                                         \
        NGP_PROF(IO_ == 1 ? "k_ffmlp_fwd_tm_sigma" : (IO_ == 3 ? "k_ffmlp_fwd_tm_sdf_loss" : "k_ffmlp_fwd_tm"), st);                                                  \
        k_ffmlp_fwd_tm<WT_, HID_, G_, IO_><<<grid_t, 128 * npt, smem_t, st>>>((const __half *)inputs, (const WT_ *)weights, (__half *)outputs, m, ntiles, cols_t, accw, pcols, batch_limit(), io); \
    }
#define NGP_FWDTM_CASE(WT_, HID_) { if (m.act) NGP_FWDTM_LAUNCH(WT_, HID_, true, 0) else NGP_FWDTM_LAUNCH(WT_, HID_, false, 0) }
#define NGP_FWDTM_HID(WT_)                                       \
    switch (hid) {                                               \
        case 16: NGP_FWDTM_CASE(WT_, 16) break;                  \
        case 32: NGP_FWDTM_CASE(WT_, 32) break;                  \
        default: NGP_FWDTM_CASE(WT_, 64) break;                  \
    }
            if (io_mode == 1) NGP_FWDTM_LAUNCH(float, 64, false, 1)
            else if (io_mode == 3) NGP_FWDTM_LAUNCH(float, 64, false, 3)
            else if (wdtype == NGP_DTYPE_F16) NGP_FWDTM_HID(__half) else NGP_FWDTM_HID(float)
#undef NGP_FWDTM_HID
#undef NGP_FWDTM_CASE
#undef NGP_FWDTM_LAUNCH
            NGP_CHECK_LAUNCH();
            return NGP_OK;
        }
    }
    if (io_mode == 3) { set_error("%s: the fused SDF loss exists for the tensor-memory forward only (option ffmlp_fwd_tmem = 1)", fn); return NGP_EUNSUPPORTED; }
    if (io_mode) {  // one tile layout, one width: two instantiations
        rc = io_mode == 1 ? set_smem(k_ffmlp_fwd<float, 64, false, 1>, smem) : set_smem(k_ffmlp_fwd<float, 64, false, 2>, smem);
        if (rc != NGP_OK) return rc;
        NGP_PROF(io_mode == 1 ? "k_ffmlp_fwd_sigma" : "k_ffmlp_fwd_color", st);
        if (io_mode == 1) k_ffmlp_fwd<float, 64, false, 1><<<grid, 128 * np, smem, st>>>((const __half *)inputs, (const float *)weights, (__half *)outputs, m, ntiles, cols, accw, batch_limit(), io);
        else k_ffmlp_fwd<float, 64, false, 2><<<grid, 128 * np, smem, st>>>((const __half *)inputs, (const float *)weights, (__half *)outputs, m, ntiles, cols, accw, batch_limit(), io);
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
#define NGP_FWD_CASE(WT_, HID_)                                                                                              \
    {                                                                                                                        \
        rc = m.act ? set_smem(k_ffmlp_fwd<WT_, HID_, true, 0>, smem) : set_smem(k_ffmlp_fwd<WT_, HID_, false, 0>, smem);         \
        if (rc != NGP_OK) return rc;                                                                                         \
        NGP_PROF("k_ffmlp_fwd", st);                                                                                         \
        if (m.act) k_ffmlp_fwd<WT_, HID_, true, 0><<<grid, 128 * np, smem, st>>>((const __half *)inputs, (const WT_ *)weights, (__half *)outputs, m, ntiles, cols, accw, batch_limit(), io); \
        else k_ffmlp_fwd<WT_, HID_, false, 0><<<grid, 128 * np, smem, st>>>((const __half *)inputs, (const WT_ *)weights, (__half *)outputs, m, ntiles, cols, accw, batch_limit(), io); \
    }
#define NGP_FWD_HID(WT_)                                         \
    switch (hid) {                                               \
        case 16: NGP_FWD_CASE(WT_, 16) break;                    \
        case 32: NGP_FWD_CASE(WT_, 32) break;                    \
        case 64: NGP_FWD_CASE(WT_, 64) break;                    \
        default: NGP_FWD_CASE(WT_, 128) break;                   \
    }
    if (wdtype == NGP_DTYPE_F16) NGP_FWD_HID(__half) else NGP_FWD_HID(float)
#undef NGP_FWD_HID
#undef NGP_FWD_CASE
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

}  // namespace
}  // namespace ngp

extern "C" {

int ngp_ffmlp_forward(const void *inputs, const void *weights, int weights_dtype, uint32_t B, uint32_t input_dim, uint32_t output_dim,
                      uint32_t hidden_dim, uint32_t num_layers, uint32_t activation, uint32_t output_activation, void *forward_buffer,
                      void *outputs, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    // hidden_dim <= 128: never written (the backward recomputes the activations); 256: the reference's [num_layers, B, 256] layout
    return ngp::run_forward("ffmlp_forward", inputs, weights, weights_dtype, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                            output_activation, forward_buffer, outputs, stream);
}

int ngp_ffmlp_inference(const void *inputs, const void *weights, int weights_dtype, uint32_t B, uint32_t input_dim, uint32_t output_dim,
                        uint32_t hidden_dim, uint32_t num_layers, uint32_t activation, uint32_t output_activation, void *inference_buffer,
                        void *outputs, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    (void)inference_buffer;
    return ngp::run_forward("ffmlp_inference", inputs, weights, weights_dtype, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                            output_activation, nullptr, outputs, stream);
}

size_t ngp_ffmlp_backward_workspace_bytes(uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers) {
    const size_t nw = (size_t)hidden_dim * (input_dim + (size_t)hidden_dim * (num_layers - 1) + output_dim);
    return (((nw + 3) & ~(size_t)3) + 4) * sizeof(float);
}

}  // extern "C"

namespace ngp {
namespace {
int run_backward(const char *fn, const void *grad, const void *inputs, const void *weights, int weights_dtype, const void *forward_buffer, uint32_t B,
                 uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                 uint32_t output_activation, int calc_grad_inputs, void *backward_buffer, void *grad_inputs, void *grad_weights,
                 int grad_weights_dtype, void *workspace, size_t workspace_bytes, void *stream, int io_mode = 0, const nerf_io_t *iop = nullptr) {
    const nerf_io_t io = iop ? *iop : nerf_io_t{};
    NGP_CHECK_ARG(grad_weights, "%s: null pointer argument", fn);
    NGP_CHECK_ARG(grad_weights_dtype == NGP_DTYPE_F16 || grad_weights_dtype == NGP_DTYPE_F32, "%s: grad_weights must be a half or float tensor", fn);
    int rc = check_dims(fn, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, weights_dtype);
    if (rc != NGP_OK) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    mlp_t m;
    make_mlp(m, B, input_dim, output_dim, hidden_dim, num_layers);
    m.act = activation;
    if (activation == 2) { set_error("%s: the Sine activation has no backward: its derivative needs the pre-activation, which is not kept (the reference's warp_activation_backward leaves it unimplemented too, utils.h:552-556)", fn); return NGP_EUNSUPPORTED; }
    if (B == 0) {  // empty batch: zero gradient
        NGP_CHECK_CUDA(cudaMemsetAsync(grad_weights, 0, (size_t)m.nweights * (grad_weights_dtype == NGP_DTYPE_F32 ? 4 : 2), st));
        return NGP_OK;
    }
    NGP_CHECK_ARG((grad || io_mode == 2) && inputs && weights && workspace, "%s: null pointer argument", fn);
    NGP_CHECK_ARG(!calc_grad_inputs || grad_inputs, "%s: calc_grad_inputs needs grad_inputs", fn);
    rc = check_io(fn, io_mode, input_dim, output_dim, hidden_dim, activation, weights_dtype);
    if (rc != NGP_OK) return rc;
    NGP_CHECK_ARG(workspace_bytes >= ngp_ffmlp_backward_workspace_bytes(input_dim, output_dim, hidden_dim, num_layers),
                  "%s: workspace too small (%zu bytes)", fn, workspace_bytes);
    if (hidden_dim == 256)
        return wide_backward(fn, grad, inputs, weights, weights_dtype, forward_buffer, B, input_dim, output_dim, num_layers, activation, calc_grad_inputs,
                             backward_buffer, grad_inputs, grad_weights, grad_weights_dtype == NGP_DTYPE_F32, (float *)workspace, st);
    bwd_cols_t tc;
    make_bwd_cols(m, calc_grad_inputs != 0, tc);
    if (tc.total > 512) { set_error("%s: configuration needs %u TMEM columns (> 512)", fn, tc.total); return NGP_EUNSUPPORTED; }
    uint32_t nbuf = 2;
    size_t smem = m.wbytes + nbuf * (size_t)(ktiles(input_dim) + ktiles(output_dim)) * kActTileBytes +
                  (2 + (size_t)num_layers) * ktiles(hidden_dim) * kActTileBytes;
    if (smem + 1024 > kMaxSmem) {  // drop the prefetch ring before giving up
        nbuf = 1;
        smem -= (size_t)(ktiles(input_dim) + ktiles(output_dim)) * kActTileBytes;
    }
    if (smem + 1024 > kMaxSmem) { set_error("%s: configuration needs %zu bytes of shared memory (> %zu)", fn, smem, kMaxSmem); return NGP_EUNSUPPORTED; }
    const uint32_t cols = tmem_cols_pow2(tc.total);
    const uint32_t ntiles = div_round_up(B, kTileRows);
    uint32_t per_sm = (uint32_t)(kMaxSmem / (smem + 1024));
    if (per_sm > 512 / cols) per_sm = 512 / cols;
    if (per_sm < 1) per_sm = 1;
    uint32_t grid = (uint32_t)sm_count() * per_sm;
    if (grid > ntiles) grid = ntiles;
    const int gf32 = grad_weights_dtype == NGP_DTYPE_F32;
    // ---- multi-pipeline variant: NP tiles in flight per CTA, one CTA per SM ----
    {
        const int variant = io_mode ? 1 : get_option("ffmlp_bwd_variant", -1);  // the fused nets exist in the multi-pipeline kernel only
        const bool packed = input_dim + output_dim <= 64;
        const size_t per_pipe = ((size_t)ktiles(input_dim) + (packed ? 0 : ktiles(output_dim)) + (size_t)num_layers * ktiles(hidden_dim)) * kActTileBytes;
        uint32_t accw = hidden_dim > input_dim ? hidden_dim : input_dim;
        accw = (accw + 31u) & ~31u;
        uint32_t dwcols = 0;
        for (uint32_t i = 0; i <= num_layers; ++i) dwcols += (m.K[i] + 31u) & ~31u;
        uint32_t np = 0;
        if (m.wbytes + 2048 < kMaxSmem) np = (uint32_t)((kMaxSmem - 2048 - m.wbytes) / per_pipe);  // 2 KB: static shared + alignment
        if (dwcols < 512) { if (np > (512 - dwcols) / accw) np = (512 - dwcols) / accw; } else np = 0;
        if (np > 4) np = 4;
        // operand columns in tensor memory for the K = 64 layer products (see the kernel): hidden 64, ReLU, self-issuing pipelines.
        // ffmlp_bwd_tmem: 0 off, 1 when it costs no pipeline, 2 even if the 32 extra columns per pipeline cost one
        const int nis = get_option("ffmlp_bwd_issuers", 3);  // 1 one central issuer, 2 two (layer products / weight gradients), 3 pipelines issue their own layer products
        const int tm_opt = get_option("ffmlp_bwd_tmem", 0);
        uint32_t opc = 0;
        if (hidden_dim == 64 && activation == 0 && nis >= 3 && tm_opt > 0 && dwcols < 512) {
            const uint32_t np_t = (512 - dwcols) / (accw + 32u);
            if (np_t >= np || (tm_opt >= 2 && np_t >= 2)) {
                opc = 32;
                if (np > np_t) np = np_t;
            }
        }
        // ns threads share a row's epilogue (latency); 64 + 128 * np * ns <= 832 threads keeps 72 registers per thread
        uint32_t ns = hidden_dim >= 64 ? 2u : 1u;
        const int ns_opt = get_option("ffmlp_bwd_split", 0);
        if (ns_opt == 1 || ns_opt == 2) ns = (uint32_t)ns_opt;
        if (hidden_dim / ns < 32 && hidden_dim >= 32) ns = hidden_dim / 32;
        if (hidden_dim < 32) ns = 1;
        // tiles in flight beat threads per tile: when one thread per row lets a fourth pipeline fit (shared memory and TMEM
        // permitting), take it (measured, sigma net 32->64x2->16 at 590 k rows: 4 x 128 threads 89 us, 3 x 256 threads 98 us)
        if (ns_opt == 0 && ns == 2 && np >= 4) ns = 1;
        if (np * ns > 6) np = 6 / ns;
        const uint32_t per_cta_need = div_round_up(ntiles, (uint32_t)sm_count());
        if (variant != 1 && np > per_cta_need) np = per_cta_need;  // forced (tests): as many pipelines as fit, whatever the batch
        if (np > ntiles) np = ntiles;
        if (variant != 0 && (np >= 2 || (variant == 1 && np >= 1))) {
            const size_t smem_mp = m.wbytes + np * per_pipe;
            int dual = (nis != 1 ? 1 : 0) | (nis >= 3 ? 2 : 0);
            const uint32_t pw = accw + opc;  // TMEM columns per pipeline
#ifdef NGP_FFMLP_DEBUG  // developer builds only: timing ablations that produce WRONG results are not reachable from the shipped library
            dual |= get_option("ffmlp_ablate", 0) << 4;
#endif
            const uint32_t cols_mp = tmem_cols_pow2(np * pw + dwcols);
            uint32_t grid_mp = div_round_up(ntiles, np);
            if (grid_mp > (uint32_t)sm_count()) grid_mp = (uint32_t)sm_count();
            if (io_mode == 2) {
                rc = set_smem(k_ffmlp_bwd_mp<float, 64, false, 2>, smem_mp);
                if (rc != NGP_OK) return rc;
                NGP_PROF("k_ffmlp_bwd_mp_color", st);
                k_ffmlp_bwd_mp<float, 64, false, 2><<<grid_mp, 64 + 128 * np * ns, smem_mp, st>>>((const __half *)grad, (const __half *)inputs, (const float *)weights, (__half *)grad_inputs,
                                                        (float *)workspace, grad_weights, gf32, m, ntiles, cols_mp, calc_grad_inputs, np, ns, pw, dual, batch_limit(), io, opc);
                NGP_CHECK_LAUNCH();
                return NGP_OK;
            }
#define NGP_BWDMP_CASE(WT_, HID_)                                                                                            \
    {                                                                                                                        \
        rc = m.act ? set_smem(k_ffmlp_bwd_mp<WT_, HID_, true, 0>, smem_mp) : set_smem(k_ffmlp_bwd_mp<WT_, HID_, false, 0>, smem_mp);   \
        if (rc != NGP_OK) return rc;                                                                                         \
        NGP_PROF("k_ffmlp_bwd_mp", st);                                                                                      \
        if (m.act) k_ffmlp_bwd_mp<WT_, HID_, true, 0><<<grid_mp, 64 + 128 * np * ns, smem_mp, st>>>((const __half *)grad, (const __half *)inputs, (const WT_ *)weights, (__half *)grad_inputs, \
                                                        (float *)workspace, grad_weights, gf32, m, ntiles, cols_mp, calc_grad_inputs, np, ns, pw, dual, batch_limit(), io, opc);   \
        else k_ffmlp_bwd_mp<WT_, HID_, false, 0><<<grid_mp, 64 + 128 * np * ns, smem_mp, st>>>((const __half *)grad, (const __half *)inputs, (const WT_ *)weights, (__half *)grad_inputs, \
                                                        (float *)workspace, grad_weights, gf32, m, ntiles, cols_mp, calc_grad_inputs, np, ns, pw, dual, batch_limit(), io, opc);   \
    }
#define NGP_BWDMP_HID(WT_)                                         \
    switch (hidden_dim) {                                          \
        case 16: NGP_BWDMP_CASE(WT_, 16) break;                    \
        case 32: NGP_BWDMP_CASE(WT_, 32) break;                    \
        case 64: NGP_BWDMP_CASE(WT_, 64) break;                    \
        default: NGP_BWDMP_CASE(WT_, 128) break;                   \
    }
            if (weights_dtype == NGP_DTYPE_F16) NGP_BWDMP_HID(__half) else NGP_BWDMP_HID(float)
#undef NGP_BWDMP_HID
#undef NGP_BWDMP_CASE
            NGP_CHECK_LAUNCH();
            return NGP_OK;
        }
    }
#define NGP_BWD_CASE(WT_, HID_)                                                                                              \
    {                                                                                                                        \
        rc = set_smem(k_ffmlp_bwd<WT_, HID_, true>, smem);                                                                         \
        if (rc != NGP_OK) return rc;                                                                                         \
        NGP_PROF("k_ffmlp_bwd", st);                                                                                         \
        k_ffmlp_bwd<WT_, HID_, true><<<grid, 128 * split_of<HID_>::NS, smem, st>>>((const __half *)grad, (const __half *)inputs, (const WT_ *)weights, (__half *)grad_inputs, \
                                                        (float *)workspace, grad_weights, gf32, m, ntiles, cols, calc_grad_inputs, nbuf, batch_limit());   \
    }
#define NGP_BWD_HID(WT_)                                         \
    switch (hidden_dim) {                                        \
        case 16: NGP_BWD_CASE(WT_, 16) break;                    \
        case 32: NGP_BWD_CASE(WT_, 32) break;                    \
        case 64: NGP_BWD_CASE(WT_, 64) break;                    \
        default: NGP_BWD_CASE(WT_, 128) break;                   \
    }
    if (weights_dtype == NGP_DTYPE_F16) NGP_BWD_HID(__half) else NGP_BWD_HID(float)
#undef NGP_BWD_HID
#undef NGP_BWD_CASE
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}
}  // namespace
}  // namespace ngp

extern "C" {

int ngp_ffmlp_backward(const void *grad, const void *inputs, const void *weights, int weights_dtype, const void *forward_buffer, uint32_t B,
                       uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                       uint32_t output_activation, int calc_grad_inputs, void *backward_buffer, void *grad_inputs, void *grad_weights,
                       int grad_weights_dtype, void *workspace, size_t workspace_bytes, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    return ngp::run_backward("ffmlp_backward", grad, inputs, weights, weights_dtype, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers,
                             activation, output_activation, calc_grad_inputs, backward_buffer, grad_inputs, grad_weights, grad_weights_dtype, workspace,
                             workspace_bytes, stream);
}

/* ---- network_ff nets with their glue fused in (include/ngp_b200.h, "next" rows) ---- */
int ngp_nerf_sigma_forward(const void *enc, const void *weights, int weights_dtype, uint32_t M, uint32_t num_layers, float density_scale, void *h,
                           float *sigmas, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    if (M == 0) return NGP_OK;
    if (!(h && sigmas)) { ngp::set_error("nerf_sigma_forward: null pointer argument"); return NGP_EINVAL; }
    ngp::nerf_io_t io{};
    io.sigmas = sigmas;
    io.density_scale = density_scale;
    return ngp::run_forward("nerf_sigma_forward", enc, weights, weights_dtype, M, 32, 16, 64, num_layers, 0, 6, nullptr, h, stream, 1, &io);
}

int ngp_sdf_forward_loss(const void *enc, const void *weights, int weights_dtype, const float *target, uint32_t B, uint32_t num_layers,
                         const float *loss_scale, void *grad_h, float *loss_out, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    if (B == 0) return NGP_OK;
    if (!(target && grad_h && loss_out)) { ngp::set_error("sdf_forward_loss: null pointer argument"); return NGP_EINVAL; }
    ngp::nerf_io_t io{};
    io.target = target;
    io.loss_scale = loss_scale;
    io.loss_out = loss_out;
    return ngp::run_forward("sdf_forward_loss", enc, weights, weights_dtype, B, 32, 16, 64, num_layers, 0, 6, nullptr, grad_h, stream, 3, &io);
}

int ngp_nerf_color_forward(const float *dirs, const void *h, const void *weights, int weights_dtype, uint32_t M, uint32_t num_layers, float *rgbs,
                           void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    if (M == 0) return NGP_OK;
    if (!(dirs && rgbs)) { ngp::set_error("nerf_color_forward: null pointer argument"); return NGP_EINVAL; }
    ngp::nerf_io_t io{};
    io.dirs = dirs;
    io.rgbs = rgbs;
    return ngp::run_forward("nerf_color_forward", h, weights, weights_dtype, M, 32, 16, 64, num_layers, 0, 6, nullptr, nullptr, stream, 2, &io);
}

int ngp_nerf_nets_forward(const void *enc, const float *dirs, const void *weights_sigma, const void *weights_color, int weights_dtype, uint32_t M,
                          uint32_t num_layers_sigma, uint32_t num_layers_color, float density_scale, void *h, float *sigmas, float *rgbs, void *stream,
                          const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    const char *fn = "nerf_nets_forward";
    if (M == 0) return NGP_OK;
    NGP_CHECK_ARG(enc && dirs && weights_sigma && weights_color && sigmas && rgbs, "%s: null pointer argument", fn);
    if (weights_dtype != NGP_DTYPE_F32 || num_layers_sigma != 2 || num_layers_color < 2 || num_layers_color > 6) {
        set_error("%s: fp32 parameters, a 2-layer sigma net and a 2..6-layer colour net (network_ff.py:26-49 uses 2 and 3)", fn);
        return NGP_EUNSUPPORTED;
    }
    mlp_t ms, mc;
    make_mlp(ms, M, 32, 16, 64, num_layers_sigma);
    make_mlp(mc, M, 32, 16, 64, num_layers_color);
    const uint32_t ntiles = div_round_up(M, kTileRows);
    const size_t per_pipe = 2 * kActTileBytes;
    uint32_t np = (uint32_t)((kMaxSmem - 1024 - ms.wbytes - mc.wbytes) / per_pipe);
    if (np > 8) np = 8;
    const uint32_t per_cta_need = div_round_up(ntiles, (uint32_t)sm_count());
    if (np > per_cta_need) np = per_cta_need;
    const size_t smem = ms.wbytes + mc.wbytes + np * per_pipe;
    const uint32_t cols = tmem_cols_pow2(np * 64u);
    uint32_t grid = div_round_up(ntiles, np);
    if (grid > (uint32_t)sm_count()) grid = (uint32_t)sm_count();
    nerf_io_t io{};
    io.dirs = dirs; io.sigmas = sigmas; io.rgbs = rgbs; io.density_scale = density_scale;
    cudaStream_t st = (cudaStream_t)stream;
    if (get_option("nerf_nets_tmem", 1)) {  // activations in tensor memory (default); 0: activation tiles in shared memory
        uint32_t npt = 5;  // 96 TMEM columns per pipeline
        if (npt > per_cta_need) npt = per_cta_need;
        const size_t smem_t = ms.wbytes + mc.wbytes;
        uint32_t grid_t = div_round_up(ntiles, npt);
        if (grid_t > (uint32_t)sm_count()) grid_t = (uint32_t)sm_count();
        int rc_t = h ? set_smem(k_nerf_nets_fwd_tm<true>, smem_t) : set_smem(k_nerf_nets_fwd_tm<false>, smem_t);
        if (rc_t != NGP_OK) return rc_t;
        NGP_PROF("k_nerf_nets_fwd_tm", st);
        if (h) k_nerf_nets_fwd_tm<true><<<grid_t, 128 * npt, smem_t, st>>>((const __half *)enc, (const float *)weights_sigma, (const float *)weights_color, (__half *)h, ms, mc, ntiles, 512u, batch_limit(), io);
        else k_nerf_nets_fwd_tm<false><<<grid_t, 128 * npt, smem_t, st>>>((const __half *)enc, (const float *)weights_sigma, (const float *)weights_color, nullptr, ms, mc, ntiles, 512u, batch_limit(), io);
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    int rc = h ? set_smem(k_nerf_nets_fwd<true>, smem) : set_smem(k_nerf_nets_fwd<false>, smem);
    if (rc != NGP_OK) return rc;
    NGP_PROF("k_nerf_nets_fwd", st);
    if (h) k_nerf_nets_fwd<true><<<grid, 128 * np, smem, st>>>((const __half *)enc, (const float *)weights_sigma, (const float *)weights_color, (__half *)h, ms, mc, ntiles, cols, batch_limit(), io);
    else k_nerf_nets_fwd<false><<<grid, 128 * np, smem, st>>>((const __half *)enc, (const float *)weights_sigma, (const float *)weights_color, nullptr, ms, mc, ntiles, cols, batch_limit(), io);
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

int ngp_nerf_color_backward(const float *grad_rgbs, const float *rgbs, const float *grad_sigmas, const float *dirs, const void *h, const void *weights,
                            int weights_dtype, uint32_t M, uint32_t num_layers, float density_scale, void *grad_h, void *grad_weights,
                            int grad_weights_dtype, void *workspace, size_t workspace_bytes, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    if (!(grad_rgbs && rgbs && grad_sigmas && dirs && grad_h) && M) { ngp::set_error("nerf_color_backward: null pointer argument"); return NGP_EINVAL; }
    ngp::nerf_io_t io{};
    io.dirs = dirs;
    io.rgbs = const_cast<float *>(rgbs);
    io.grad_rgbs = grad_rgbs;
    io.grad_sigmas = grad_sigmas;
    io.density_scale = density_scale;
    return ngp::run_backward("nerf_color_backward", nullptr, h, weights, weights_dtype, nullptr, M, 32, 16, 64, num_layers, 0, 6, 1, nullptr, grad_h,
                             grad_weights, grad_weights_dtype, workspace, workspace_bytes, stream, 2, &io);
}

}  // extern "C"

#ifdef NGP_FFMLP_TRACE
extern "C" __attribute__((visibility("default"))) int ngp_ffmlp_trace_dump(unsigned long long *events, unsigned *counts, int reset) {
    cudaDeviceSynchronize();
    if (events) cudaMemcpyFromSymbol(events, ngp::g_tr, sizeof(unsigned long long) * 8 * 1024);
    if (counts) cudaMemcpyFromSymbol(counts, ngp::g_trn, sizeof(unsigned) * 8);
    if (reset) { unsigned z[8] = {0, 0, 0, 0, 0, 0, 0, 0}; cudaMemcpyToSymbol(ngp::g_trn, z, sizeof(z)); }
    return 0;
}
#endif
