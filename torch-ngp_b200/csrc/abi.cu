// abi.cu -- error reporting, options and version of the C ABI (include/ngp_b200.h).
#include <stdarg.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace ngp {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static std::mutex g_opt_mu;
static std::map<std::string, int> &options() {
    static std::map<std::string, int> o = {
        {"hashgrid_fwd_variant", -1},  // -1 auto, 0 generic, 1 paired-gather (point,level) threads, 2 sector-gather (point, 4 levels) threads, 3 = 2 with per-level-kind index code
        {"hashgrid_bwd_variant", -1},  // -1 auto, 0 generic atomics, 1 paired + warp-aggregated (+ smem levels), 2 (point, 4 levels) threads
        {"hashgrid_v3_minb_fwd", -1},  // v3 forward: resident blocks per SM the register allocation targets (4 or 6; -1 = by dtype)
        {"hashgrid_v3_minb_bwd", -1},  // v3 backward: same
        {"hashgrid_v3_mix_fwd", 0},    // v3 forward: warps of one block work on different level groups (1) or whole blocks do (0)
        {"hashgrid_v3_mix_bwd", 1},    // v3 backward: same choice (measured: helps the backward on ray-ordered samples, hurts the forward on random points)
        {"hashgrid_v3_unroll_bwd", -1}, // v3 backward: four levels of a thread unrolled (1) or one loop body (0: 4x smaller code, no instruction-cache thrash: the default)
        {"hashgrid_bwd_red4", 2},      // v3 backward, fp16 table: one red.v4.f16x2 per 4-entry group holding both x-neighbours: 0 never, 1 always, 2 when the lanes of a warp do not share groups (random points)
#ifdef NGP_FFMLP_DEBUG
        {"ffmlp_ablate", 0},           // developer builds only: timing experiments with WRONG results (1 skip the weight-gradient MMAs, 2 the layer MMAs, 3 both)
#endif
        {"march_rays_warp", 32768},    // inference marcher: one warp per alive ray while n_alive <= this (0 = always one thread per ray, the reference's mapping)
        {"hashgrid_smem_levels", 1},   // keep coarse levels in shared memory (fast paths)
        {"hashgrid_bwd_warp_agg", 1},  // collapse warp-uniform cells before the L2 reduction
        {"march_variant", -1},         // training marcher: -1 auto / 1 warp per ray (ballot replay of the skip rule), 0 thread per ray
        {"march_masks", 1},            // warp marcher: record emit masks in pass 1 when the workspace has room (0: always re-march)
        {"composite_variant", -1},     // training compositor: -1 auto / 1 warp per ray (scans), 0 thread per ray
        {"ffmlp_bwd_tmem", 0},         // FFMLP backward, hidden 64: K = 64 layer products read A from tensor memory: 0 off (default: measured no gain, the backward is bound by its per-round chain, not by operand traffic), 1 when it costs no tile pipeline, 2 always
        {"ffmlp_fwd_tmem", 1},         // FFMLP forward, hidden <= 64: activations in tensor memory (1) or in shared-memory tiles (0); same results
        {"nerf_nets_tmem", 1},         // ngp_nerf_nets_forward: activations in tensor memory, A operand read from TMEM (1) or activation tiles in shared memory (0); same results
        {"ffmlp_available", 1},
        {"ffmlp_bwd_variant", -1},     // -1 auto / 1: multi-pipeline backward (one MMA-issuer warp, up to 4 tiles in flight), 0: one tile per CTA
        {"ffmlp_bwd_issuers", 3},      // 3: every pipeline issues its own layer products, warp 1 the weight gradients; 2: warp 0 / warp 1; 1: a single issuer warp
        {"ffmlp_bwd_split", 0},        // 0 auto / 1 / 2: threads sharing one accumulator row in the multi-pipeline backward epilogue
        {"ffmlp_fwd_pipes", 0},        // > 0: cap on the tile pipelines per CTA of the FFMLP forward (tuning)
        {"profile_events", 0},         // 1: bracket every kernel launch with CUDA events (ngp_profile_collect)        // read-only in spirit: 1 once the tcgen05 FFMLP kernels are linked in
    };
    return o;
}

int get_option(const char *key, int dflt) {
    std::lock_guard<std::mutex> l(g_opt_mu);
    auto it = options().find(key);
    return it == options().end() ? dflt : it->second;
}

// ---- launch accounting ---------------------------------------------------------------------------
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
struct prof_rec { const char *name; cudaEvent_t a, b; };
static std::mutex g_prof_mu;
static std::vector<prof_rec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_free;

static cudaEvent_t prof_event() {
    if (!g_prof_free.empty()) { cudaEvent_t e = g_prof_free.back(); g_prof_free.pop_back(); return e; }
    cudaEvent_t e = nullptr;
    cudaEventCreate(&e);
    return e;
}

prof_scope::prof_scope(const char *name, cudaStream_t st_) : slot(-1), st(st_) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> l(g_prof_mu);
    prof_rec r{name, prof_event(), prof_event()};
    cudaEventRecord(r.a, st);
    slot = (int)g_prof_recs.size();
    g_prof_recs.push_back(r);
}
prof_scope::~prof_scope() {
    if (slot < 0) return;
    std::lock_guard<std::mutex> l(g_prof_mu);
    cudaEventRecord(g_prof_recs[slot].b, st);
}

// Per-call options (include/ngp_b200.h ngp_call_opts): an extern "C" entry point opens an opts_scope for the duration
// of the call; the launch code below it reads the accessors.  Thread-local, so nothing is shared between callers.
static thread_local const ngp_call_opts *t_opts = nullptr;
opts_scope::opts_scope(const ngp_call_opts *o) : prev(t_opts) { t_opts = o; }
opts_scope::~opts_scope() { t_opts = prev; }
const int *batch_limit() { return t_opts ? reinterpret_cast<const int *>(t_opts->batch_limit) : nullptr; }
float input_bound() { return (t_opts && t_opts->input_bound > 0.f) ? t_opts->input_bound : 0.f; }
uint32_t call_flags() { return t_opts ? t_opts->flags : 0u; }
int grid_cap_blocks_per_sm() { return t_opts ? t_opts->grid_cap_blocks_per_sm : 0; }
void level_range(uint32_t &begin, uint32_t &end) { begin = t_opts ? t_opts->level_begin : 0u; end = t_opts ? t_opts->level_end : 0u; }

int sm_count() {
    static int cached[64] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

}  // namespace ngp

extern "C" {

const char *ngp_last_error(void) { return ngp::g_err; }
int ngp_abi_version(void) { return 2; }  // 2: ngp_call_opts grew level_begin / level_end

int ngp_set_option(const char *key, int value) {
    if (!key) return NGP_EINVAL;
    std::lock_guard<std::mutex> l(ngp::g_opt_mu);
    auto it = ngp::options().find(key);
    if (it == ngp::options().end()) {
        ngp::set_error("unknown option '%s'", key);
        return NGP_EINVAL;
    }
    it->second = value;
    if (strcmp(key, "profile_events") == 0) ngp::g_prof_on.store(value != 0);
    return NGP_OK;
}

long long ngp_launch_count(void) { return ngp::g_launches.load(); }

// Synchronises the recorded events and writes one line per kernel name: "<name> <count> <total_ms>\n".
// Returns the number of bytes written (excluding the NUL), or a negative error code; clears the records.
int ngp_profile_collect(char *buf, size_t buf_bytes) {
    using namespace ngp;
    if (!buf || buf_bytes == 0) return NGP_EINVAL;
    std::lock_guard<std::mutex> l(g_prof_mu);
    std::map<std::string, std::pair<long long, double>> agg;
    for (auto &r : g_prof_recs) {
        float ms = 0.f;
        if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
            auto &e = agg[r.name];
            e.first += 1;
            e.second += ms;
        }
        g_prof_free.push_back(r.a);
        g_prof_free.push_back(r.b);
    }
    g_prof_recs.clear();
    cudaGetLastError();
    size_t off = 0;
    buf[0] = 0;
    for (auto &kv : agg) {
        int n = snprintf(buf + off, buf_bytes - off, "%s %lld %.6f\n", kv.first.c_str(), kv.second.first, kv.second.second);
        if (n < 0 || (size_t)n >= buf_bytes - off) break;
        off += (size_t)n;
    }
    return (int)off;
}

int ngp_get_option(const char *key, int *value) {
    if (!key || !value) return NGP_EINVAL;
    std::lock_guard<std::mutex> l(ngp::g_opt_mu);
    auto it = ngp::options().find(key);
    if (it == ngp::options().end()) {
        ngp::set_error("unknown option '%s'", key);
        return NGP_EINVAL;
    }
    *value = it->second;
    return NGP_OK;
}

int ngp_allocate_splitk(size_t) { return NGP_OK; }
int ngp_free_splitk(void) { return NGP_OK; }

}  // extern "C"
