// hashgrid_host.cuh -- host-side dispatch + C ABI entry points of the grid encoder.
// Included by hashgrid.cu only.
#pragma once

namespace ngp {
namespace {

constexpr size_t kSmemLevelBudget = 200 * 1024;  // bytes of a level slice we are willing to stage per block

template <typename T, uint32_t D>
int launch_fwd_generic(const float *inputs, const T *emb, const int *offsets, T *out, uint32_t B, uint32_t C, uint32_t L,
                       float S, uint32_t H, T *dy_dx, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st) {
    const dim3 grid(div_round_up(B, 256u), L, 1);
    const bool quirk = (call_flags() & NGP_OPT_EXACT_INPUT_GRAD) == 0;  // default: the reference's dy_dx (gridencoder.cu:143)
    switch (C) {
        case 1: { NGP_PROF("k_grid_fwd_generic", st); k_grid_fwd_generic<T, D, 1><<<grid, 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, dy_dx, gridtype, ac, interp, layout, quirk, batch_limit()); } break;
        case 2: { NGP_PROF("k_grid_fwd_generic", st); k_grid_fwd_generic<T, D, 2><<<grid, 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, dy_dx, gridtype, ac, interp, layout, quirk, batch_limit()); } break;
        case 4: { NGP_PROF("k_grid_fwd_generic", st); k_grid_fwd_generic<T, D, 4><<<grid, 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, dy_dx, gridtype, ac, interp, layout, quirk, batch_limit()); } break;
        case 8: { NGP_PROF("k_grid_fwd_generic", st); k_grid_fwd_generic<T, D, 8><<<grid, 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, dy_dx, gridtype, ac, interp, layout, quirk, batch_limit()); } break;
        default: set_error("GridEncoding: C must be 1, 2, 4, or 8."); return NGP_EINVAL;
    }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

template <typename T, uint32_t D>
int launch_bwd_generic(const T *grad, const float *inputs, const int *offsets, T *g_emb, uint32_t B, uint32_t C, uint32_t L,
                       float S, uint32_t H, const T *dy_dx, T *g_in, uint32_t gridtype, bool ac, uint32_t interp, int layout,
                       cudaStream_t st) {
    const uint32_t N_C = C < 2 ? C : 2;
    const dim3 grid(div_round_up(B * C / N_C, 256u), L, 1);
    const uint32_t gi = div_round_up(B * D, 256u);
#define NGP_BWD_CASE(CC)                                                                                                  \
    case CC:                                                                                                              \
        { NGP_PROF("k_grid_bwd_generic", st); k_grid_bwd_generic<T, D, CC><<<grid, 256, 0, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac, interp, layout, batch_limit()); } \
        if (dy_dx) { NGP_PROF("k_input_bwd", st); k_input_bwd<T, D, CC><<<gi, 256, 0, st>>>(grad, dy_dx, g_in, B, L, layout); }                            \
        break;
    switch (C) {
        NGP_BWD_CASE(1) NGP_BWD_CASE(2) NGP_BWD_CASE(4) NGP_BWD_CASE(8)
        default: set_error("GridEncoding: C must be 1, 2, 4, or 8."); return NGP_EINVAL;
    }
#undef NGP_BWD_CASE
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

template <typename T, uint32_t D>
int launch_tv(const T *inputs, const T *emb, T *grad, const int *offsets, float weight, uint32_t B, uint32_t C, uint32_t L,
              float S, uint32_t H, uint32_t gridtype, bool ac, cudaStream_t st) {
    const dim3 grid(div_round_up(B, 256u), L, 1);
    switch (C) {
        case 1: { NGP_PROF("k_grad_tv", st); k_grad_tv<T, D, 1><<<grid, 256, 0, st>>>(inputs, emb, grad, offsets, weight, B, L, S, H, gridtype, ac); } break;
        case 2: { NGP_PROF("k_grad_tv", st); k_grad_tv<T, D, 2><<<grid, 256, 0, st>>>(inputs, emb, grad, offsets, weight, B, L, S, H, gridtype, ac); } break;
        case 4: { NGP_PROF("k_grad_tv", st); k_grad_tv<T, D, 4><<<grid, 256, 0, st>>>(inputs, emb, grad, offsets, weight, B, L, S, H, gridtype, ac); } break;
        case 8: { NGP_PROF("k_grad_tv", st); k_grad_tv<T, D, 8><<<grid, 256, 0, st>>>(inputs, emb, grad, offsets, weight, B, L, S, H, gridtype, ac); } break;
        default: set_error("GridEncoding: C must be 1, 2, 4, or 8."); return NGP_EINVAL;
    }
    NGP_CHECK_LAUNCH();
    return NGP_OK;
}

#define NGP_SWITCH_D(D_, CALL)                                                     \
    switch (D_) {                                                                  \
        case 2: { constexpr uint32_t kD = 2; return CALL; }                        \
        case 3: { constexpr uint32_t kD = 3; return CALL; }                        \
        case 4: { constexpr uint32_t kD = 4; return CALL; }                        \
        case 5: { constexpr uint32_t kD = 5; return CALL; }                        \
        default: set_error("GridEncoding: D must be 2, 3, 4, or 5."); return NGP_EINVAL; \
    }

// The fast path needs to know the level slices on the host (alignment of the wide accesses,
// which levels fit shared memory).  offsets_host == NULL -> generic kernels.
bool fast_path_ok(const int *offsets_host, uint32_t L) {
    if (!offsets_host) return false;
    for (uint32_t l = 0; l <= L; ++l)
        if (offsets_host[l] & 7) return false;
    return true;
}

// number of leading levels served from shared memory
template <typename T>
uint32_t plan_smem_levels(const int *offsets_host, uint32_t L, uint32_t B, size_t bytes_per_entry, size_t *max_bytes) {
    *max_bytes = 0;
    if (!get_option("hashgrid_smem_levels", 1) || B < 32768) return 0;
    uint32_t n = 0;
    for (; n < L; ++n) {
        const size_t bytes = (size_t)(offsets_host[n + 1] - offsets_host[n]) * bytes_per_entry;
        if (bytes > kSmemLevelBudget) break;
        if (bytes > *max_bytes) *max_bytes = bytes;
    }
    return n;
}

template <typename T>
int fwd_d3c2(const float *inputs, const T *emb, const int *offsets, const int *offsets_host, T *out, uint32_t B, uint32_t L,
             float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st) {
    const int variant = get_option("hashgrid_fwd_variant", -1);
    if ((variant == -1 || variant == 3) && L % 4 == 0 && ((uintptr_t)emb & 31) == 0 && ((uintptr_t)out & 31) == 0) {
        const uint32_t ng = L / 4;
        const bool mix = get_option("hashgrid_v3_mix_fwd", 0) && (ng == 1 || ng == 2 || ng == 4 || ng == 8);
        const dim3 gr = mix ? dim3(div_round_up(B, 256u / ng), 1, 1) : dim3(div_round_up(B, 256u), ng, 1);
        const int minb = get_option("hashgrid_v3_minb_fwd", -1);
        const bool occ6 = minb < 0 ? sizeof(T) == 2 : minb >= 6;  // fp16 groups are 4 registers, fp32 groups 8: 40 registers spill the latter
        NGP_PROF("k_grid_fwd_v3", st);
#define NGP_FWD_V3(MINB, MIX) k_grid_fwd_v3<T, MINB, MIX><<<gr, 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, gridtype, ac, interp, layout, input_bound(), batch_limit())
        if (mix) { if (occ6) NGP_FWD_V3(6, true); else NGP_FWD_V3(4, true); }
        else { if (occ6) NGP_FWD_V3(6, false); else NGP_FWD_V3(4, false); }
#undef NGP_FWD_V3
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    if (input_bound() > 0.f) { set_error("grid_encode_forward: ngp_call_opts.input_bound needs the v3 kernels (L %% 4 == 0, 32-byte aligned buffers)"); return NGP_EUNSUPPORTED; }
    if (variant == 2 && L % 4 == 0 && ((uintptr_t)emb & 31) == 0 && ((uintptr_t)out & 31) == 0) {
        { NGP_PROF("k_grid_fwd_d3c2_lg4", st); k_grid_fwd_d3c2_lg4<T><<<dim3(div_round_up(B, 256u), L / 4, 1), 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, gridtype, ac, interp, layout, batch_limit()); }
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    size_t smem = 0;
    const uint32_t ns = plan_smem_levels<T>(offsets_host, L, B, 2 * sizeof(T), &smem);
    if (ns > 0) {
        NGP_CHECK_CUDA(cudaFuncSetAttribute(k_grid_fwd_d3c2_smem<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLevelBudget));

        const uint32_t per_sm = smem <= 100 * 1024 ? 2 : 1;
        uint32_t nb = (uint32_t)sm_count() * per_sm / ns;
        nb = nb < 1 ? 1 : nb;
        const uint32_t need = div_round_up(B, 1024u);
        nb = nb > need ? need : nb;
        { NGP_PROF("k_grid_fwd_d3c2_smem", st); k_grid_fwd_d3c2_smem<T><<<dim3(nb, ns, 1), 1024, smem, st>>>(inputs, emb, offsets, out, B, L, S, H, gridtype, ac, interp, layout, batch_limit()); }
        NGP_CHECK_LAUNCH();
    }
    if (ns < L) {
        { NGP_PROF("k_grid_fwd_d3c2", st); k_grid_fwd_d3c2<T><<<dim3(div_round_up(B, 256u), L - ns, 1), 256, 0, st>>>(inputs, emb, offsets, out, B, L, S, H, gridtype, ac,
                                                                                 interp, layout, ns, batch_limit()); }
        NGP_CHECK_LAUNCH();
    }
    return NGP_OK;
}

template <typename T>
int bwd_d3c2(const T *grad, const float *inputs, const int *offsets, const int *offsets_host, T *g_emb, uint32_t B, uint32_t L,
             float S, uint32_t H, uint32_t gridtype, bool ac, uint32_t interp, int layout, cudaStream_t st) {
    const int variant = get_option("hashgrid_bwd_variant", -1);
    if ((variant == -1 || variant == 3) && L % 4 == 0 && ((uintptr_t)g_emb & 31) == 0 && ((uintptr_t)grad & 31) == 0) {
        const int warp_agg = get_option("hashgrid_bwd_warp_agg", 1);
        const uint32_t ng = L / 4;
        // ngp_call_opts.level_begin / level_end: this call accumulates a range of level groups only (blockIdx.y-major launch)
        uint32_t lb = 0, le = 0;
        level_range(lb, le);
        const bool ranged = le > 0;
        if (ranged && (lb % 4 || le % 4 || lb >= le || le > L)) { set_error("grid_encode_backward: level range [%u, %u) must be multiples of 4 inside [0, %u]", lb, le, L); return NGP_EINVAL; }
        const uint32_t ngc = ranged ? (le - lb) / 4 : ng, lg_first = ranged ? lb / 4 : 0;  // level groups of this launch
        const bool mix = get_option("hashgrid_v3_mix_bwd", 1) && (ngc == 1 || ngc == 2 || ngc == 4 || ngc == 8);
        const dim3 gr = mix ? dim3(div_round_up(B, 256u / ngc), 1, 1) : dim3(div_round_up(B, 256u), ngc, 1);
        const int minb = get_option("hashgrid_v3_minb_bwd", -1);
        const bool occ6 = minb < 0 ? false : minb >= 6;
        const int red4 = (call_flags() & NGP_OPT_GRID_BWD_GROUP_RED) ? 1 : get_option("hashgrid_bwd_red4", 2);
        // Unordered points (NGP_OPT_GRID_BWD_SMEM_COARSE; the SDF step, configs[0]): every point lands in one of a few
        // thousand coarse cells and the L2 atomic unit serialises same-address reductions.  The levels that fit are
        // accumulated in shared memory instead (fp32, one flush per block).  Ray-ordered samples keep them in the warp
        // reduction (runs of equal cells collapse before they reach L2).
        // Shared-memory atomics are no faster per operation than the L2 unit (ATOMS: 2 cycles per lane and SM, i.e. ~140 G/s
        // over 148 SMs; an exact fixed-point variant on the native integer ATOMS.ADD with levels cut into slabs measured
        // 204 us for levels 0-1 against 101 + 92 us here): what pays is FEWER reductions, which is what ngp_bin_points +
        // the warp reduction give the SDF step (nerf/sdf_trainer.py bin_bits).
        uint32_t skip = 0;
        if ((call_flags() & NGP_OPT_GRID_BWD_SMEM_COARSE) && lb == 0) {
            size_t smem = 0;
            const uint32_t ns = plan_smem_levels<T>(offsets_host, L, B, 2 * sizeof(float), &smem);
            if (ns > 0 && (!ranged || ns <= le)) {
                NGP_CHECK_CUDA(cudaFuncSetAttribute(k_grid_bwd_d3c2_smem<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLevelBudget));

                const uint32_t per_sm = smem <= 100 * 1024 ? 2 : 1;
                uint32_t nb = (uint32_t)sm_count() * per_sm / ns;
                nb = nb < 1 ? 1 : nb;
                const uint32_t need = div_round_up(B, 1024u);
                nb = nb > need ? need : nb;
                { NGP_PROF("k_grid_bwd_d3c2_smem", st); k_grid_bwd_d3c2_smem<T><<<dim3(nb, ns, 1), 1024, smem, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac, interp, layout, batch_limit(), input_bound()); }
                NGP_CHECK_LAUNCH();
                skip = ns;
            }
        }
        NGP_PROF("k_grid_bwd_v3", st);
        const int un_opt = get_option("hashgrid_v3_unroll_bwd", -1);
        const bool unroll = un_opt > 0;  // default: one loop body (A/B in one run: -13 % per ray-ordered sample, no difference on unordered points)
#define NGP_BWD_V3(MINB, MIX) { if (unroll) k_grid_bwd_v3<T, MINB, MIX, true><<<gr, 256, 0, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac, interp, layout, warp_agg, red4, skip, input_bound(), batch_limit(), lg_first, ngc); \
                                else k_grid_bwd_v3<T, MINB, MIX, false><<<gr, 256, 0, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac, interp, layout, warp_agg, red4, skip, input_bound(), batch_limit(), lg_first, ngc); }
        if (mix) { if (occ6) NGP_BWD_V3(6, true) else NGP_BWD_V3(4, true) }
        else { if (occ6) NGP_BWD_V3(6, false) else NGP_BWD_V3(4, false) }
#undef NGP_BWD_V3
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    { uint32_t lb = 0, le = 0; level_range(lb, le);
      if (le > 0) { set_error("grid_encode_backward: ngp_call_opts.level_end needs the v3 kernels (L %% 4 == 0, 32-byte aligned buffers)"); return NGP_EUNSUPPORTED; } }
    if (input_bound() > 0.f) { set_error("grid_encode_backward: ngp_call_opts.input_bound needs the v3 kernels (L %% 4 == 0, 32-byte aligned buffers)"); return NGP_EUNSUPPORTED; }
    if (variant == 2 && L % 4 == 0 && ((uintptr_t)g_emb & 31) == 0 && ((uintptr_t)grad & 31) == 0) {
        const int warp_agg = get_option("hashgrid_bwd_warp_agg", 1);
        { NGP_PROF("k_grid_bwd_d3c2_lg4", st); k_grid_bwd_d3c2_lg4<T><<<dim3(div_round_up(B, 256u), L / 4, 1), 256, 0, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac, interp, layout, warp_agg, batch_limit()); }
        NGP_CHECK_LAUNCH();
        return NGP_OK;
    }
    size_t smem = 0;
    const uint32_t ns = plan_smem_levels<T>(offsets_host, L, B, 2 * sizeof(float), &smem);  // fp32 accumulators
    if (ns > 0) {
        NGP_CHECK_CUDA(cudaFuncSetAttribute(k_grid_bwd_d3c2_smem<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemLevelBudget));

        const uint32_t per_sm = smem <= 100 * 1024 ? 2 : 1;
        uint32_t nb = (uint32_t)sm_count() * per_sm / ns;
        nb = nb < 1 ? 1 : nb;
        const uint32_t need = div_round_up(B, 1024u);
        nb = nb > need ? need : nb;
        { NGP_PROF("k_grid_bwd_d3c2_smem", st); k_grid_bwd_d3c2_smem<T><<<dim3(nb, ns, 1), 1024, smem, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac, interp, layout, batch_limit()); }
        NGP_CHECK_LAUNCH();
    }
    if (ns < L) {
        const int warp_agg = get_option("hashgrid_bwd_warp_agg", 1);
        { NGP_PROF("k_grid_bwd_d3c2", st); k_grid_bwd_d3c2<T><<<dim3(div_round_up(B, 256u), L - ns, 1), 256, 0, st>>>(grad, inputs, offsets, g_emb, B, L, S, H, gridtype, ac,
                                                                                 interp, layout, ns, warp_agg, batch_limit()); }
        NGP_CHECK_LAUNCH();
    }
    return NGP_OK;
}

template <typename T>
int fwd_typed(const float *inputs, const T *emb, const int *offsets, const int *offsets_host, T *out, uint32_t B, uint32_t D,
              uint32_t C, uint32_t L, float S, uint32_t H, T *dy_dx, uint32_t gridtype, bool ac, uint32_t interp, int layout,
              cudaStream_t st) {
    NGP_SWITCH_D(D, (launch_fwd_generic<T, kD>(inputs, emb, offsets, out, B, C, L, S, H, dy_dx, gridtype, ac, interp, layout, st)))
}
template <typename T>
int bwd_typed(const T *grad, const float *inputs, const int *offsets, T *g_emb, uint32_t B, uint32_t D, uint32_t C, uint32_t L,
              float S, uint32_t H, const T *dy_dx, T *g_in, uint32_t gridtype, bool ac, uint32_t interp, int layout,
              cudaStream_t st) {
    NGP_SWITCH_D(D, (launch_bwd_generic<T, kD>(grad, inputs, offsets, g_emb, B, C, L, S, H, dy_dx, g_in, gridtype, ac, interp, layout, st)))
}
template <typename T>
int tv_typed(const T *inputs, const T *emb, T *grad, const int *offsets, float weight, uint32_t B, uint32_t D, uint32_t C,
             uint32_t L, float S, uint32_t H, uint32_t gridtype, bool ac, cudaStream_t st) {
    NGP_SWITCH_D(D, (launch_tv<T, kD>(inputs, emb, grad, offsets, weight, B, C, L, S, H, gridtype, ac, st)))
}

}  // namespace
}  // namespace ngp

extern "C" {

int ngp_grid_encode_forward(const float *inputs, const void *embeddings, const int *offsets, const int *offsets_host,
                            void *outputs, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, void *dy_dx,
                            uint32_t gridtype, int align_corners, uint32_t interp, int dtype, int layout, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    if (B == 0) return NGP_OK;
    NGP_CHECK_ARG(inputs && embeddings && offsets && outputs, "grid_encode_forward: null pointer argument");
    NGP_CHECK_ARG(layout == NGP_LAYOUT_LBC || layout == NGP_LAYOUT_BLC, "grid_encode_forward: bad layout %d", layout);
    NGP_CHECK_ARG(L >= 1 && L <= 65535, "grid_encode_forward: L=%u out of range", L);
    if (B == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool ac = align_corners != 0;
    const int variant = get_option("hashgrid_fwd_variant", -1);
    const bool fast = D == 3 && C == 2 && !dy_dx && variant != 0 && fast_path_ok(offsets_host, L);
    if (input_bound() > 0.f && !(fast && (dtype == NGP_DTYPE_F32 || dtype == NGP_DTYPE_F16))) {
        set_error("grid_encode_forward: ngp_call_opts.input_bound is only supported by the D=3, C=2 fast kernels (no dy_dx, offsets_host given)");
        return NGP_EUNSUPPORTED;
    }
    switch (dtype) {
        case NGP_DTYPE_F32:
            if (fast) return fwd_d3c2<float>(inputs, (const float *)embeddings, offsets, offsets_host, (float *)outputs, B, L, S, H, gridtype, ac, interp, layout, st);
            return fwd_typed<float>(inputs, (const float *)embeddings, offsets, offsets_host, (float *)outputs, B, D, C, L, S, H, (float *)dy_dx, gridtype, ac, interp, layout, st);
        case NGP_DTYPE_F16:
            if (fast) return fwd_d3c2<__half>(inputs, (const __half *)embeddings, offsets, offsets_host, (__half *)outputs, B, L, S, H, gridtype, ac, interp, layout, st);
            return fwd_typed<__half>(inputs, (const __half *)embeddings, offsets, offsets_host, (__half *)outputs, B, D, C, L, S, H, (__half *)dy_dx, gridtype, ac, interp, layout, st);
        case NGP_DTYPE_F64:
            return fwd_typed<double>(inputs, (const double *)embeddings, offsets, offsets_host, (double *)outputs, B, D, C, L, S, H, (double *)dy_dx, gridtype, ac, interp, layout, st);
        default: set_error("grid_encode_forward: embeddings must be a floating tensor (dtype code %d)", dtype); return NGP_EINVAL;
    }
}

int ngp_grid_encode_backward(const void *grad, const float *inputs, const void *embeddings, const int *offsets,
                             const int *offsets_host, void *grad_embeddings, uint32_t B, uint32_t D, uint32_t C, uint32_t L,
                             float S, uint32_t H, const void *dy_dx, void *grad_inputs, uint32_t gridtype, int align_corners,
                             uint32_t interp, int dtype, int layout, void *stream, const ngp_call_opts *opts) {
    ngp::opts_scope _os(opts);
    using namespace ngp;
    (void)embeddings;
    if (B == 0) return NGP_OK;
    NGP_CHECK_ARG(grad && inputs && offsets && grad_embeddings, "grid_encode_backward: null pointer argument");
    NGP_CHECK_ARG((dy_dx == nullptr) == (grad_inputs == nullptr), "grid_encode_backward: dy_dx and grad_inputs must be given together");
    NGP_CHECK_ARG(layout == NGP_LAYOUT_LBC || layout == NGP_LAYOUT_BLC, "grid_encode_backward: bad layout %d", layout);
    NGP_CHECK_ARG(L >= 1 && L <= 65535, "grid_encode_backward: L=%u out of range", L);
    if (B == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool ac = align_corners != 0;
    const int variant = get_option("hashgrid_bwd_variant", -1);
    const bool fast = D == 3 && C == 2 && variant != 0 && fast_path_ok(offsets_host, L);
    if (input_bound() > 0.f && !(fast && !dy_dx && (dtype == NGP_DTYPE_F32 || dtype == NGP_DTYPE_F16))) {
        set_error("grid_encode_backward: ngp_call_opts.input_bound is only supported by the D=3, C=2 fast kernels (no dy_dx, offsets_host given)");
        return NGP_EUNSUPPORTED;
    }
    {
        uint32_t lb = 0, le = 0;
        level_range(lb, le);
        if (le > 0 && !(fast && !dy_dx && (dtype == NGP_DTYPE_F32 || dtype == NGP_DTYPE_F16))) {
            set_error("grid_encode_backward: ngp_call_opts.level_end is only supported by the D=3, C=2 fast kernels (no dy_dx, offsets_host given)");
            return NGP_EUNSUPPORTED;
        }
    }
    int rc;
    switch (dtype) {
        case NGP_DTYPE_F32:
            if (fast) {
                rc = bwd_d3c2<float>((const float *)grad, inputs, offsets, offsets_host, (float *)grad_embeddings, B, L, S, H, gridtype, ac, interp, layout, st);
                if (rc != NGP_OK || !dy_dx) return rc;
                { NGP_PROF("k_input_bwd", st); k_input_bwd<float, 3, 2><<<div_round_up(B * 3, 256u), 256, 0, st>>>((const float *)grad, (const float *)dy_dx, (float *)grad_inputs, B, L, layout); }
                NGP_CHECK_LAUNCH();
                return NGP_OK;
            }
            return bwd_typed<float>((const float *)grad, inputs, offsets, (float *)grad_embeddings, B, D, C, L, S, H, (const float *)dy_dx, (float *)grad_inputs, gridtype, ac, interp, layout, st);
        case NGP_DTYPE_F16:
            if (fast) {
                rc = bwd_d3c2<__half>((const __half *)grad, inputs, offsets, offsets_host, (__half *)grad_embeddings, B, L, S, H, gridtype, ac, interp, layout, st);
                if (rc != NGP_OK || !dy_dx) return rc;
                { NGP_PROF("k_input_bwd", st); k_input_bwd<__half, 3, 2><<<div_round_up(B * 3, 256u), 256, 0, st>>>((const __half *)grad, (const __half *)dy_dx, (__half *)grad_inputs, B, L, layout); }
                NGP_CHECK_LAUNCH();
                return NGP_OK;
            }
            return bwd_typed<__half>((const __half *)grad, inputs, offsets, (__half *)grad_embeddings, B, D, C, L, S, H, (const __half *)dy_dx, (__half *)grad_inputs, gridtype, ac, interp, layout, st);
        case NGP_DTYPE_F64:
            return bwd_typed<double>((const double *)grad, inputs, offsets, (double *)grad_embeddings, B, D, C, L, S, H, (const double *)dy_dx, (double *)grad_inputs, gridtype, ac, interp, layout, st);
        default: set_error("grid_encode_backward: grad must be a floating tensor (dtype code %d)", dtype); return NGP_EINVAL;
    }
}

int ngp_grad_total_variation(const void *inputs, const void *embeddings, void *grad, const int *offsets, float weight,
                             uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, uint32_t gridtype,
                             int align_corners, int dtype, void *stream) {
    using namespace ngp;
    if (B == 0) return NGP_OK;
    NGP_CHECK_ARG(inputs && embeddings && grad && offsets, "grad_total_variation: null pointer argument");
    if (B == 0) return NGP_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const bool ac = align_corners != 0;
    switch (dtype) {
        case NGP_DTYPE_F32: return tv_typed<float>((const float *)inputs, (const float *)embeddings, (float *)grad, offsets, weight, B, D, C, L, S, H, gridtype, ac, st);
        case NGP_DTYPE_F16: return tv_typed<__half>((const __half *)inputs, (const __half *)embeddings, (__half *)grad, offsets, weight, B, D, C, L, S, H, gridtype, ac, st);
        case NGP_DTYPE_F64: return tv_typed<double>((const double *)inputs, (const double *)embeddings, (double *)grad, offsets, weight, B, D, C, L, S, H, gridtype, ac, st);
        default: set_error("grad_total_variation: embeddings must be a floating tensor (dtype code %d)", dtype); return NGP_EINVAL;
    }
}

}  // extern "C"
