// tmem_ld_bench.cu -- how fast can the warps of one SM read TMEM (tcgen05.ld 32x32b.x32) and write shared memory?  Informs the
// FFMLP drain budget.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o benchmarks/_build/tmem_ld_bench benchmarks/tmem_ld_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../torch-ngp_b200/csrc/umma.cuh"
using namespace ngp::umma;

// nwg warpgroups (128 threads each); every thread drains `reps` x 64 columns of its TMEM lane (2 x ld.x32), optionally packs to
// fp16 and stores 128 B to a swizzled shared-memory tile (mode 1), like an FFMLP epilogue without the tensor core.
__global__ void __launch_bounds__(1024) k_ld(uint32_t reps, int mode, long long *out, float *sink) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    if (threadIdx.x < 32) tmem_alloc(&s_tmem, 512);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t warp = threadIdx.x >> 5, wg = threadIdx.x >> 7, row = threadIdx.x & 127u;
    const uint32_t trow = s_tmem + (((warp & 3u) * 32u) << 16) + (wg & 7u) * 64u;
    uint8_t *tile = smem + wg * 16384u + row * 128u;
    float acc = 0.f;
    __syncthreads();
    const long long t0 = clock64();
    for (uint32_t r = 0; r < reps; ++r) {
        for (uint32_t c0 = 0; c0 < 64; c0 += 32) {
            float v[32];
            tmem_ld32(trow + c0, v);
            if (mode == 0) {
#pragma unroll
                for (int j = 0; j < 32; ++j) acc += v[j];
            } else {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                    uint4 u;
                    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(u.x) : "f"(v[j + 1]), "f"(v[j]));
                    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(u.y) : "f"(v[j + 3]), "f"(v[j + 2]));
                    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(u.z) : "f"(v[j + 5]), "f"(v[j + 4]));
                    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(u.w) : "f"(v[j + 7]), "f"(v[j + 6]));
                    *reinterpret_cast<uint4 *>(tile + ((((c0 >> 3) + (j >> 3)) ^ (row & 7u)) << 4)) = u;
                }
            }
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) out[0] = t1 - t0;
    if (acc == 123.456f) sink[0] = acc;
    fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(s_tmem, 512);
}

int main() {
    long long *out; float *sink;
    cudaMallocManaged(&out, 16); cudaMalloc(&sink, 16);
    cudaFuncSetAttribute(k_ld, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const uint32_t reps = 256;
    for (int mode = 0; mode < 2; ++mode)
        for (uint32_t nwg : {1u, 2u, 4u, 6u, 8u}) {
            for (int rep = 0; rep < 2; ++rep) k_ld<<<1, 128 * nwg, 200 * 1024>>>(reps, mode, out, sink);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            const double bytes = (double)reps * nwg * 128 * 64 * 4;  // fp32 accumulator bytes drained
            printf("mode %d (%s) warpgroups %u: %.0f cycles, %.1f B/clk of TMEM read, %.0f cycles per 128x64 drain per warpgroup\n", mode,
                   mode ? "ld + relu/pack + st.shared" : "ld + add", nwg, (double)out[0], bytes / out[0], (double)out[0] / reps);
        }
    return 0;
}
