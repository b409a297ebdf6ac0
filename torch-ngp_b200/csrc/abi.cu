// abi.cu -- error reporting, options and version of the C ABI (include/ngp_b200.h).
#include <stdarg.h>
#include <string.h>

#include <map>
#include <mutex>
#include <string>

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
        {"hashgrid_fwd_variant", -1},  // -1 auto, 0 generic, 1 paired-gather fast path
        {"hashgrid_bwd_variant", -1},  // -1 auto, 0 generic atomics, 1 paired + warp-aggregated
        {"hashgrid_smem_levels", 1},   // keep coarse levels in shared memory (fast paths)
        {"hashgrid_bwd_warp_agg", 1},  // collapse warp-uniform cells before the L2 reduction
        {"hashgrid_ref_pos_deriv_quirk", 0},  // 1: reproduce the reference's zero dy_dx for d>0 (linear interp)
        {"march_block", 128},
    };
    return o;
}

int get_option(const char *key, int dflt) {
    std::lock_guard<std::mutex> l(g_opt_mu);
    auto it = options().find(key);
    return it == options().end() ? dflt : it->second;
}

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
int ngp_abi_version(void) { return 1; }

int ngp_set_option(const char *key, int value) {
    if (!key) return NGP_EINVAL;
    std::lock_guard<std::mutex> l(ngp::g_opt_mu);
    auto it = ngp::options().find(key);
    if (it == ngp::options().end()) {
        ngp::set_error("unknown option '%s'", key);
        return NGP_EINVAL;
    }
    it->second = value;
    return NGP_OK;
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
