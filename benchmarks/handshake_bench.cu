// handshake_bench.cu -- cost of one epilogue <-> MMA-issuer hand-off round on B200, piece by piece.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o benchmarks/_build/handshake_bench benchmarks/handshake_bench.cu
// One issuer warp + NP epilogue warpgroups (128 threads).  Round: epilogue arrives on full[p] -> issuer sees it ->
// (optionally) issues NMMA tcgen05.mma -> signals done[p] (tcgen05.commit or a plain arrive) -> epilogue wakes up.
// flags: 1 = tcgen05.commit (else mbarrier.arrive), 2 = fence.proxy.async in the epilogue, 4 = tcgen05 fences,
//        8 = epilogue reads 64 TMEM columns (tcgen05.ld) per round, 16 = issuer polls all pipelines with one ballot,
//        32 = one lane per epilogue warp polls the done barrier
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../torch-ngp_b200/csrc/umma.cuh"
using namespace ngp::umma;

__global__ void __launch_bounds__(32 + 128 * 4) k_hs(uint32_t np, uint32_t rounds, uint32_t nmma, uint32_t flags, long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_full[4], s_done[4];
    for (uint32_t i = threadIdx.x; i < 32 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) tmem_alloc(&s_tmem, 512);
    if (threadIdx.x < np) { mbar_init(&s_full[threadIdx.x], 4); mbar_init(&s_done[threadIdx.x], 1); }
    if (threadIdx.x == 0) mbar_init_fence();
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t idesc = make_idesc(128, 64, false, false);
    const uint64_t ad = make_desc(smem_u32(smem), 0, kAtomBytes), bd = make_desc(smem_u32(smem + 16384), 0, kAtomBytes);
    const long long t0 = clock64();
    if (warp == 0) {
        uint32_t left[4], ph[4], active = np;
        for (uint32_t p = 0; p < 4; ++p) { left[p] = p < np ? rounds : 0; ph[p] = 0; }
        if (flags & 16u) {
            // one poll for all pipelines: lane p tests full[p], one ballot tells the warp which are ready
            uint32_t phbits = 0, myleft = lane < np ? rounds : 0;
            while (__any_sync(0xffffffffu, myleft != 0)) {
                const bool ready = myleft != 0 && mbar_test(&s_full[lane & 3u], (phbits >> (lane & 3u)) & 1u);
                const uint32_t m = __ballot_sync(0xffffffffu, ready);
                if (m == 0) continue;
                phbits ^= m;
                if (ready) --myleft;
                if (flags & 4u) fence_after_sync();
#pragma unroll
                for (uint32_t p = 0; p < 4; ++p) {
                    if (!(m & (1u << p))) continue;
                    if (lane == 0) {
                        for (uint32_t k = 0; k < nmma; ++k) mma_f16(tmem + p * 64u, ad, bd, idesc, k > 0 ? 1u : 0u);
                        if (flags & 1u) mma_commit(&s_done[p]); else mbar_arrive(&s_done[p]);
                    }
                }
                __syncwarp();
            }
            active = 0;
        }
        while (active) {
#pragma unroll
            for (uint32_t p = 0; p < 4; ++p) {
                const bool go = p < np && left[p] && mbar_test(&s_full[p], ph[p]);
                if (!__any_sync(0xffffffffu, go)) continue;
                ph[p] ^= 1u;
                if (flags & 4u) fence_after_sync();
                if (lane == 0) {
                    for (uint32_t k = 0; k < nmma; ++k) mma_f16(tmem + p * 64u, ad, bd, idesc, k > 0 ? 1u : 0u);
                    if (flags & 1u) mma_commit(&s_done[p]); else mbar_arrive(&s_done[p]);
                }
                __syncwarp();
                if (--left[p] == 0) --active;
            }
        }
    } else {
        const uint32_t p = (warp - 1) >> 2, q = warp & 3u;
        if (p < np) {
            uint32_t phase = 0;
            float acc = 0.f;
            for (uint32_t r = 0; r < rounds; ++r) {
                if (flags & 2u) fence_proxy_async();
                if (flags & 4u) fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(&s_full[p]);
                if (flags & 32u) {  // one polling lane per warp
                    if (lane == 0) mbar_wait(&s_done[p], phase);
                    __syncwarp();
                } else {
                    mbar_wait(&s_done[p], phase);
                }
                phase ^= 1u;
                if (flags & 4u) fence_after_sync();
                if (flags & 8u) {
                    float v[64];
                    tmem_ld64(tmem + p * 64u + ((q * 32u) << 16), v);
                    acc += v[0] + v[63];
                }
            }
            if (acc == 123.f) out[7] = 1;
        }
    }
    fence_before_sync();
    __syncthreads();
    if (threadIdx.x == 0) out[0] = clock64() - t0;
    if (warp == 0) tmem_dealloc(tmem, 512);
}

int main() {
    long long *out;
    cudaMallocManaged(&out, 64);
    cudaFuncSetAttribute(k_hs, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const uint32_t rounds = 2000;
    for (uint32_t np : {1u, 3u, 4u})
        for (uint32_t nmma : {0u, 4u, 12u})
            for (uint32_t flags : {1u, 17u, 31u, 49u, 63u}) {
                if (nmma > 0 && !(flags & 1u)) continue;  // MMAs need the commit to be waited on
                for (int rep = 0; rep < 2; ++rep) {
                    k_hs<<<1, 32 + 128 * np, 64 * 1024>>>(np, rounds, nmma, flags, out);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                }
                printf("np=%u nmma=%2u flags=%2u (%s%s%s%s%s%s): %.0f cycles per round of one pipeline, %.0f per round over all pipelines\n", np, nmma, flags,
                       flags & 1 ? "commit " : "arrive ", flags & 2 ? "proxyfence " : "", flags & 4 ? "tcfences " : "", flags & 8 ? "tmemld " : "", flags & 16 ? "ballotpoll " : "", flags & 32 ? "lanewait" : "",
                       (double)out[0] / rounds, (double)out[0] / rounds / np);
            }
    return 0;
}
