// common.cuh -- shared host/device helpers of libngp_b200.so (sm_100a only).
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/ngp_b200.h"

namespace ngp {

void set_error(const char *fmt, ...);
int get_option(const char *key, int dflt);
int sm_count();  // SMs of the current device (148 on B200), cached
// Per-call options (ngp_call_opts, include/ngp_b200.h): valid inside an opts_scope opened by the extern "C" entry point
// on the calling thread; defaults (nullptr / 0) outside.  No process-wide state.
struct opts_scope {
    const ngp_call_opts *prev;
    explicit opts_scope(const ngp_call_opts *o);
    ~opts_scope();
};
const int *batch_limit();        // opts->batch_limit (device int) or nullptr
float input_bound();             // opts->input_bound if > 0, else 0: the grid kernels map [-bound, bound] -> [0, 1] themselves
uint32_t call_flags();           // opts->flags (NGP_OPT_*)
int grid_cap_blocks_per_sm();    // opts->grid_cap_blocks_per_sm
void level_range(uint32_t &begin, uint32_t &end);  // opts->level_begin / level_end (0, 0 = every level)

// Effective batch of a per-sample kernel: min(B, *limit * mult) when a device-side limit is set.  The launch is
// still sized for B; blocks past the limit exit at once.  `limit` was written by an earlier kernel on the stream.
__device__ __forceinline__ uint32_t eff_batch(uint32_t B, const int *__restrict__ limit, uint32_t mult = 1u) {
    if (limit == nullptr) return B;
    const int v = __ldg(limit);
    const unsigned long long lim = (unsigned long long)(v > 0 ? v : 0) * mult;
    return lim < B ? (uint32_t)lim : B;
}

// Per-launch accounting (abi.cu).  Always counts launches (ngp_launch_count); when the option
// "profile_events" is 1 it also brackets the launch with CUDA events on `st`, read back through
// ngp_profile_collect() -- bench.py's live per-kernel durations.  Not capture-safe: keep it off
// while a CUDA graph is being captured.
struct prof_scope {
    int slot;
    cudaStream_t st;
    prof_scope(const char *name, cudaStream_t st);
    ~prof_scope();
};
#define NGP_PROF_CAT2(a, b) a##b
#define NGP_PROF_CAT(a, b) NGP_PROF_CAT2(a, b)
#define NGP_PROF(name, st) ngp::prof_scope NGP_PROF_CAT(_ngp_ps_, __LINE__)(name, st)

template <typename T>
__host__ __device__ inline T div_round_up(T a, T b) { return (a + b - 1) / b; }

#define NGP_CHECK_ARG(cond, ...)                 \
    do {                                         \
        if (!(cond)) {                           \
            ngp::set_error(__VA_ARGS__);         \
            return NGP_EINVAL;                   \
        }                                        \
    } while (0)

#define NGP_CHECK_CUDA(expr)                                                                  \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess) {                                                              \
            ngp::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return NGP_ECUDA;                                                                 \
        }                                                                                     \
    } while (0)

// Launch-error check: cudaPeekAtLastError is capture-safe and does not synchronise.
#define NGP_CHECK_LAUNCH()                                                                    \
    do {                                                                                      \
        cudaError_t _e = cudaPeekAtLastError();                                               \
        if (_e != cudaSuccess) {                                                              \
            ngp::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
            return NGP_ECUDA;                                                                 \
        }                                                                                     \
    } while (0)

}  // namespace ngp
