// umma_selftest.cu -- diagnostics entry point: runs a caller-described sequence of tcgen05.mma
// instructions on caller-provided shared-memory byte images and dumps the TMEM accumulator.
// tests/test_gpu_umma.py uses it to pin the operand-image / descriptor conventions of umma.cuh
// (K-major and MN-major views of the 128B-swizzled tile) against a plain fp32 matmul.
#include "common.cuh"
#include "umma.cuh"

namespace ngp {
namespace {

__global__ void __launch_bounds__(128) k_umma_raw(const uint8_t *__restrict__ a_img, uint32_t a_bytes, const uint8_t *__restrict__ b_img,
                                                  uint32_t b_bytes, float *__restrict__ D, uint32_t N, uint32_t nk, uint32_t idesc,
                                                  uint32_t a_lbo, uint32_t a_sbo, uint32_t a_step, uint32_t b_lbo, uint32_t b_sbo,
                                                  uint32_t b_step) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(8) uint64_t s_bar;
    uint8_t *sa = smem;
    uint8_t *sb = smem + ((a_bytes + 1023u) & ~1023u);
    for (uint32_t i = threadIdx.x; i < a_bytes / 16; i += blockDim.x) reinterpret_cast<uint4 *>(sa)[i] = reinterpret_cast<const uint4 *>(a_img)[i];
    for (uint32_t i = threadIdx.x; i < b_bytes / 16; i += blockDim.x) reinterpret_cast<uint4 *>(sb)[i] = reinterpret_cast<const uint4 *>(b_img)[i];
    if (threadIdx.x < 32) umma::tmem_alloc(&s_tmem, 256);
    if (threadIdx.x == 0) {
        umma::mbar_init(&s_bar, 1);
        umma::mbar_init_fence();
    }
    umma::fence_proxy_async();
    umma::fence_before_sync();
    __syncthreads();
    umma::fence_after_sync();
    const uint32_t tmem = s_tmem;
    if (threadIdx.x == 0) {
        for (uint32_t k = 0; k < nk; ++k) {
            const uint64_t ad = umma::make_desc(umma::smem_u32(sa) + k * a_step, a_lbo, a_sbo);
            const uint64_t bd = umma::make_desc(umma::smem_u32(sb) + k * b_step, b_lbo, b_sbo);
            umma::mma_f16(tmem, ad, bd, idesc, k > 0 ? 1u : 0u);
        }
        umma::mma_commit(&s_bar);
    }
    umma::mbar_wait(&s_bar, 0);
    umma::fence_after_sync();
    const uint32_t warp = threadIdx.x >> 5, row = threadIdx.x;
    for (uint32_t c0 = 0; c0 < N; c0 += 16) {
        float v[16];
        umma::tmem_ld16(tmem + ((warp * 32u) << 16) + c0, v);
        for (int i = 0; i < 16; ++i) D[(size_t)row * N + c0 + i] = v[i];
    }
    umma::fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) umma::tmem_dealloc(tmem, 256);
}

}  // namespace
}  // namespace ngp

extern "C" int ngp_umma_selftest(const void *a_img, uint32_t a_bytes, const void *b_img, uint32_t b_bytes, float *D, uint32_t N,
                                 uint32_t nk, uint32_t idesc, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_step, uint32_t b_lbo,
                                 uint32_t b_sbo, uint32_t b_step, void *stream) {
    using namespace ngp;
    NGP_CHECK_ARG(a_img && b_img && D, "umma_selftest: null pointer argument");
    NGP_CHECK_ARG(N % 16 == 0 && N >= 16 && N <= 256 && a_bytes % 16 == 0 && b_bytes % 16 == 0, "umma_selftest: bad N / sizes");
    const size_t smem = ((a_bytes + 1023u) & ~1023u) + ((b_bytes + 1023u) & ~1023u);
    NGP_CHECK_ARG(smem <= 200 * 1024, "umma_selftest: images too large");
    NGP_CHECK_CUDA(cudaFuncSetAttribute(k_umma_raw, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    { NGP_PROF("k_umma_raw", (cudaStream_t)stream); k_umma_raw<<<1, 128, smem, (cudaStream_t)stream>>>((const uint8_t *)a_img, a_bytes, (const uint8_t *)b_img, b_bytes, D, N, nk, idesc,
                                                       a_lbo, a_sbo, a_step, b_lbo, b_sbo, b_step); }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}
