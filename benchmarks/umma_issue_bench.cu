// umma_issue_bench.cu -- how long does one tcgen05.mma (kind::f16, M=128, K=16) take to ISSUE and to EXECUTE on B200 as a
// function of N and of how the descriptors reach the instruction?  Informs the FFMLP kernel design.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o benchmarks/_build/umma_issue_bench benchmarks/umma_issue_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../torch-ngp_b200/csrc/umma.cuh"
using namespace ngp::umma;

// mode 0: descriptors rebuilt from a runtime address per MMA (what issue_kmajor does)
// mode 1: one descriptor pair reused (pure instruction rate)
// mode 2: descriptor low words advanced by a constant (add only)
// mode 3 / 4: one descriptor pair, 4 / 2 independent TMEM accumulators in rotation
template <int MODE>
__global__ void __launch_bounds__(128) k_bench(uint32_t N, uint32_t reps, uint32_t nk, long long *out, uint32_t m_mn, uint32_t Mdim = 128) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar;
    for (uint32_t i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x < 32) tmem_alloc(&s_tmem, 512);
    if (threadIdx.x == 0) { mbar_init(&s_bar, 1); mbar_init_fence(); }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t idesc = make_idesc(Mdim, N, m_mn != 0, m_mn != 0);
    const uint32_t a0 = smem_u32(smem), b0 = smem_u32(smem + 16384);
    long long t0 = 0, t1 = 0, t2 = 0;
    uint32_t phase = 0;
    if (threadIdx.x == 0) {
        t0 = clock64();
        for (uint32_t r = 0; r < reps; ++r) {
            if (MODE == 0) {
                for (uint32_t k = 0; k < nk; ++k) {
                    const uint64_t ad = make_desc(a0 + (k >> 2) * 16384u * 0u + (k & 3u) * 32u, 0, kAtomBytes);
                    const uint64_t bd = make_desc(b0 + (k & 3u) * 32u, 0, kAtomBytes);
                    mma_f16(tmem, ad, bd, idesc, k > 0 ? 1u : 0u);
                }
            } else if (MODE == 1) {
                const uint64_t ad = make_desc(a0, 0, kAtomBytes), bd = make_desc(b0, 0, kAtomBytes);
                for (uint32_t k = 0; k < nk; ++k) mma_f16(tmem, ad, bd, idesc, 1u);
            } else if (MODE == 3) {  // round-robin over 4 independent accumulators (is the 64-cycle figure a dependency latency?)
                const uint64_t ad = make_desc(a0, 0, kAtomBytes), bd = make_desc(b0, 0, kAtomBytes);
                for (uint32_t k = 0; k < nk; ++k) mma_f16(tmem + (k & 3u) * 128u, ad, bd, idesc, 1u);
            } else if (MODE == 4) {  // 2 independent accumulators
                const uint64_t ad = make_desc(a0, 0, kAtomBytes), bd = make_desc(b0, 0, kAtomBytes);
                for (uint32_t k = 0; k < nk; ++k) mma_f16(tmem + (k & 1u) * 128u, ad, bd, idesc, 1u);
            } else {
                uint64_t ad = make_desc(a0, 0, kAtomBytes), bd = make_desc(b0, 0, kAtomBytes);
#pragma unroll 4
                for (uint32_t k = 0; k < nk; ++k) {
                    mma_f16(tmem, ad, bd, idesc, 1u);
                    ad += 2; bd += 2;
                    if ((k & 3u) == 3u) { ad -= 8; bd -= 8; }
                }
            }
        }
        t1 = clock64();
        mma_commit(&s_bar);
    }
    mbar_wait(&s_bar, phase);
    if (threadIdx.x == 0) {
        t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
    }
    fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

// several issuing warps at once (one elected lane each, own accumulator, own commit barrier): is the ~64-cycle figure a
// per-thread issue cost or the tensor pipe's?
__global__ void __launch_bounds__(128) k_multi(uint32_t N, uint32_t reps, uint32_t nk, long long *out, uint32_t niss, uint32_t m_mn = 0, uint32_t mixed = 0) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar[4];
    for (uint32_t i = threadIdx.x; i < 48 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0;
    if (threadIdx.x < 32) tmem_alloc(&s_tmem, 512);
    if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(&s_bar[i], 1); mbar_init_fence(); }
    fence_proxy_async();
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem, warp = threadIdx.x >> 5;
    // m_mn: both operands MN-major (the weight-gradient products G^T H); mixed: odd issuers MN-major, even ones K-major
    const bool mn = m_mn && (!mixed || (warp & 1u));
    const uint32_t idesc = make_idesc(128, N, mn, mn);
    const uint64_t ad = make_desc(smem_u32(smem), 0, kAtomBytes), bd = make_desc(smem_u32(smem + 16384), 0, kAtomBytes);
    const long long t0 = clock64();
    if ((threadIdx.x & 31u) == 0 && warp < niss) {
        for (uint32_t r = 0; r < reps; ++r)
            for (uint32_t k = 0; k < nk; ++k) mma_f16(tmem + warp * 128u, ad + (mn ? (uint64_t)(k & 7u) * 128u : (uint64_t)(k & 3u) * 2u), bd + (mn ? (uint64_t)(k & 7u) * 128u : (uint64_t)(k & 3u) * 2u), idesc, 1u);
        mma_commit(&s_bar[warp]);
    }
    if (warp < niss) mbar_wait(&s_bar[warp], 0);
    fence_before_sync();
    __syncthreads();
    if (threadIdx.x == 0) out[1] = clock64() - t0;
    if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

int main() {
    long long *out;
    cudaMallocManaged(&out, 16);
    const uint32_t reps = 64, nk = 8;
    for (int mn = 0; mn < 2; ++mn)
        for (uint32_t N : {16u, 32u, 64u, 128u, 256u})
            for (int mode = 0; mode < 5; ++mode) {
                if (N > 128 && mode >= 3) continue;
                auto run = [&](auto kern) {
                    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
                    kern<<<1, 128, 64 * 1024>>>(N, reps, nk, out, mn, 128u);
                    kern<<<1, 128, 64 * 1024>>>(N, reps, nk, out, mn, 128u);
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return; }
                    printf("major=%s N=%3u mode=%d  issue %.1f cyc/mma   complete %.1f cyc/mma\n", mn ? "MN" : "K ", N, mode, (double)out[0] / (reps * nk),
                           (double)out[1] / (reps * nk));
                };
                if (mode == 0) run(k_bench<0>); else if (mode == 1) run(k_bench<1>); else if (mode == 2) run(k_bench<2>); else if (mode == 3) run(k_bench<3>); else run(k_bench<4>);
            }
    // M = 64 instruction shape: is a 64-row product cheaper to issue / execute than M = 128?
    for (int mn = 0; mn < 2; ++mn)
        for (uint32_t N : {32u, 64u, 128u, 256u}) {
            cudaFuncSetAttribute(k_bench<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
            for (int rep = 0; rep < 2; ++rep) k_bench<1><<<1, 128, 64 * 1024>>>(N, reps, nk, out, mn, 64);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            printf("M=64 major=%s N=%3u  issue %.1f cyc/mma   complete %.1f cyc/mma\n", mn ? "MN" : "K ", N, (double)out[0] / (reps * nk), (double)out[1] / (reps * nk));
        }
    cudaFuncSetAttribute(k_multi, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (uint32_t N : {16u, 32u, 64u, 128u})
        for (uint32_t niss : {1u, 2u, 4u}) {
            for (int rep = 0; rep < 2; ++rep) k_multi<<<1, 128, 64 * 1024>>>(N, reps, nk, out, niss);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            printf("multi-issuer N=%3u issuers=%u  complete %.1f cyc per mma (aggregate over issuers)\n", N, niss, (double)out[1] / (reps * nk * niss));
        }
    for (uint32_t mixed : {0u, 1u})
        for (uint32_t N : {32u, 64u})
            for (uint32_t niss : {1u, 2u, 4u}) {
                for (int rep = 0; rep < 2; ++rep) k_multi<<<1, 128, 64 * 1024>>>(N, reps, nk, out, niss, 1u, mixed);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                printf("multi-issuer %s N=%3u issuers=%u  complete %.1f cyc per mma\n", mixed ? "MN/K mixed" : "MN-major  ", N, niss, (double)out[1] / (reps * nk * niss));
            }
    return 0;
}
