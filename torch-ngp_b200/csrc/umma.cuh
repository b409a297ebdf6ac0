// umma.cuh -- hand-written sm_100a tensor-core plumbing: tcgen05.mma (UMMA) with shared-memory
// operand descriptors, TMEM allocation / loads, mbarrier completion.  No CUTLASS/CuTe.
//
// Operand images.  Every fp16 operand tile lives in shared memory as a "128B-swizzled tile":
//   R rows x 64 halves (128 bytes per row), 8-row groups of 1024 bytes stored back to back,
//   the 16-byte chunk c (0..7) of row r stored at chunk position c ^ (r & 7).
// The same byte image can be consumed two ways by the tensor core:
//   K-major  : rows are the M/N index, the 64 halves are K           (SBO = 1024, LBO unused)
//   MN-major : the 64 halves are the M/N index, rows are K          (SBO = 1024 between 8-row K groups,
//                                                                     LBO = byte distance to the next 64-wide M/N slab)
// so activations written once serve both the forward product (X W^T, K-major) and the transposed
// products of the backward pass (G^T H, G W; MN-major) without any transposition.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>

namespace ngp {
namespace umma {

constexpr uint32_t kTileRowBytes = 128;   // 64 halves
constexpr uint32_t kAtomBytes = 1024;     // 8 rows

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// byte offset of 16-byte chunk `c` of row `r` inside a swizzled tile
__device__ __forceinline__ uint32_t swz_off(uint32_t r, uint32_t c) { return r * kTileRowBytes + ((c ^ (r & 7u)) << 4); }

// ---- descriptors -----------------------------------------------------------------------------
// shared-memory matrix descriptor (PTX ISA "tcgen05 shared memory descriptor"):
//   [0,14) start address >> 4 | [16,30) leading byte offset >> 4 | [32,46) stride byte offset >> 4
//   [46,48) version = 1 (sm_100) | [49,52) base offset = 0 (tiles are 1024 B aligned) | [61,64) swizzle: 2 = 128B
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor for kind::f16, fp16 x fp16 -> fp32:
//   [4,6) D format = 1 (f32) | [7,10) A = 0 (f16) | [10,13) B = 0 (f16) | 15 A major | 16 B major (0 = K, 1 = MN)
//   [17,23) N >> 3 | [24,29) M >> 4
__host__ __device__ constexpr uint32_t make_idesc(uint32_t M, uint32_t N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) | ((N >> 3) << 17) | ((M >> 4) << 24);
}

// ---- TMEM --------------------------------------------------------------------------------------
// one full warp; ncols: power of two >= 32; the allocated base address is written to *dst_smem
__device__ __forceinline__ void tmem_alloc(uint32_t *dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// generic-proxy shared-memory writes -> visible to the tensor core (async proxy)
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives row (lane base + t)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float *v) {
    uint32_t r[32];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// 32 lanes x 64 consecutive fp32 columns in one instruction (one wait)
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, float *v) {
    uint32_t r[64];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31]), "=r"(r[32]), "=r"(r[33]), "=r"(r[34]), "=r"(r[35]), "=r"(r[36]),
          "=r"(r[37]), "=r"(r[38]), "=r"(r[39]), "=r"(r[40]), "=r"(r[41]), "=r"(r[42]), "=r"(r[43]), "=r"(r[44]), "=r"(r[45]),
          "=r"(r[46]), "=r"(r[47]), "=r"(r[48]), "=r"(r[49]), "=r"(r[50]), "=r"(r[51]), "=r"(r[52]), "=r"(r[53]), "=r"(r[54]),
          "=r"(r[55]), "=r"(r[56]), "=r"(r[57]), "=r"(r[58]), "=r"(r[59]), "=r"(r[60]), "=r"(r[61]), "=r"(r[62]), "=r"(r[63])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = __uint_as_float(r[i]);
}
// n in {16, 32, 64}: one instruction, one wait
template <int N> __device__ __forceinline__ void tmem_ldN(uint32_t taddr, float *v) {
    if constexpr (N == 64) tmem_ld64(taddr, v);
    else if constexpr (N == 32) tmem_ld32(taddr, v);
    else tmem_ld16(taddr, v);
}

// 32 lanes x 16 consecutive 32-bit columns, registers -> TMEM: thread t of the warp writes row (lane base + t).  Used to put an
// fp16 A operand (two K elements per 32-bit column, row = lane) where tcgen05.mma can read it without touching shared memory.
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t *r) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
          "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// ---- MMA issue (one thread) ----------------------------------------------------------------------
// D[tmem] (+)= A[smem desc] * B[smem desc];  accumulate == 0 overwrites D
__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// the same with the A operand in TMEM ([128 lanes = rows] x [8 columns = 16 fp16 K elements] at a_tmem): only B crosses the
// shared-memory port (N x 32 B instead of 4 KB + N x 32 B per instruction)
__device__ __forceinline__ void mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t"
        "}\n" ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate), "r"(0u)
        : "memory");
}
// arrive on an mbarrier when every previously issued MMA of this thread has completed
// (implies tcgen05.fence::before_thread_sync)
__device__ __forceinline__ void mma_commit(uint64_t *mbar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(mbar)) : "memory");
}

// ---- mbarrier ------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t *mbar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(mbar)), "r"(count) : "memory");
}
// one lane of a converged warp (elect.sync).  Branching on this instead of `lane == 0` lets the compiler keep the
// tcgen05 instructions of the region in the uniform datapath without an ELECT/BRA waterfall around each of them.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t"
        ".reg .pred P;\n\t"
        "elect.sync _|P, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, P;\n\t"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void mbar_arrive(uint64_t *mbar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(mbar)) : "memory");
}
// one NON-BLOCKING probe of the phase with the given parity.  test_wait, not try_wait: try_wait may suspend the thread
// for a hardware time window when the phase is not complete -- measured on B200, a round-robin poll over four
// barriers built on try_wait cost ~1.8k cycles per hand-off.
__device__ __forceinline__ bool mbar_test(uint64_t *mbar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(mbar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// Blocking wait on ONE barrier.  The suspend-time hint matters: without it try_wait gives up after ~140 cycles and the loop
// re-issues it -- measured with ncu on the FFMLP backward, those retries were 25 % of all warp instructions of the kernel,
// taken from the issue slots of the warps that had work.  With the hint the warp stays suspended in hardware until the phase
// completes (the wake-up itself is not delayed by it).
__device__ __forceinline__ void mbar_wait(uint64_t *mbar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}\n" ::"r"(smem_u32(mbar)), "r"(parity), "r"(0x989680u)
        : "memory");
}

// ---- operand tiles ---------------------------------------------------------------------------------
// cooperative copy of a row-major fp16 matrix block [rows x cols] (cols multiple of 8, <= 64, row stride
// `ld` halves, 16 B aligned) from global memory into a swizzled tile.  All `nthreads` threads call it.
__device__ __forceinline__ void load_tile(uint8_t *tile, const __half *__restrict__ g, uint32_t rows, uint32_t cols, size_t ld,
                                          uint32_t tid, uint32_t nthreads) {
    const uint32_t cpr = cols >> 3;  // 16-byte chunks per row
    for (uint32_t q = tid; q < rows * cpr; q += nthreads) {
        const uint32_t r = q / cpr, c = q - r * cpr;
        const uint4 v = __ldg(reinterpret_cast<const uint4 *>(g + (size_t)r * ld) + c);
        *reinterpret_cast<uint4 *>(tile + swz_off(r, c)) = v;
    }
}
// the same through cp.async (LDGSTS): no register staging, completion via cp.async.wait_group
__device__ __forceinline__ void load_tile_async(uint8_t *tile, const __half *__restrict__ g, uint32_t rows, uint32_t cols, size_t ld,
                                                uint32_t tid, uint32_t nthreads) {
    const uint32_t cpr = cols >> 3;
    for (uint32_t q = tid; q < rows * cpr; q += nthreads) {
        const uint32_t r = q / cpr, c = q - r * cpr;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(tile + swz_off(r, c))),
                     "l"(reinterpret_cast<const uint4 *>(g + (size_t)r * ld) + c)
                     : "memory");
    }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Issue D[128 x N] (+)= A * B over K (multiple of 16), K-major A image (128 rows, ceil(K/64) tiles of a_tile_bytes) and
// B given either K-major ([N rows x K], tiles of b_tile_bytes) or MN-major ([K rows x N], N <= 64 per slab).
__device__ __forceinline__ void issue_kmajor(uint32_t d_tmem, uint32_t a_base, uint32_t a_tile_bytes, uint32_t b_base,
                                             uint32_t b_tile_bytes, uint32_t N, uint32_t K, uint32_t accumulate_first) {
    const uint32_t idesc = make_idesc(128, N, false, false);
    for (uint32_t k16 = 0; k16 < K / 16; ++k16) {
        const uint32_t kt = k16 >> 2, kk = k16 & 3;
        const uint64_t ad = make_desc(a_base + kt * a_tile_bytes + kk * 32, 0, kAtomBytes);
        const uint64_t bd = make_desc(b_base + kt * b_tile_bytes + kk * 32, 0, kAtomBytes);
        mma_f16(d_tmem, ad, bd, idesc, (k16 > 0 || accumulate_first) ? 1u : 0u);
    }
}

}  // namespace umma
}  // namespace ngp
