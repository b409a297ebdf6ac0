/*
 * ngp_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked into, imported by or called from the
 * product path under torch-ngp_b200/).
 *
 * A plain-C CPU restatement of the algorithms of lizuoyue/torch-ngp's Instant-NGP hot path,
 * written from the reference's CUDA sources as the specification:
 *
 *   hash grid   : /root/reference/gridencoder/src/gridencoder.cu
 *   ray marching: /root/reference/raymarching/src/raymarching.cu
 *
 * Parity status: the reference ships NO golden vectors / known-answer tests for this path
 * (SURVEY.md section 4).  This restatement is pinned instead by (i) the reference's own
 * self-checks re-run against it (fp64 gradcheck of testing/test_hashgrid_grad.py, the vertex
 * index identity of test_torch_ngp.py:159-184) and (ii) the reference's own CUDA extensions,
 * compiled unmodified by oracle/build_ref_ext.py and executed on the GPU box by tests/ on the
 * same seeded inputs (tests/test_vs_reference_ext.py).  Fixtures captured from (ii) are
 * committed under tests/golden/.
 *
 * Floating-point contract.  nvcc compiles the reference with -fmad=true, so every `a*b+c`
 * in the CUDA source is one fused multiply-add.  This file is compiled with
 * -ffp-contract=off and spells each fused operation as fmaf() explicitly, so that the
 * float sequence is the one the GPU executes (bit-exact marching needs it).
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC ngp_oracle.c -o libngp_oracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>

#define ORC_MAX_D 5
#define ORC_MAX_C 8

/* ------------------------------------------------------------------------------------------ */
/* hash grid: index                                                                            */
/* ------------------------------------------------------------------------------------------ */

/* gridencoder.cu:50-63 -- XOR of coordinate*prime, uint32 wrap-around arithmetic. */
static uint32_t orc_hash(const uint32_t *p, uint32_t D) {
    static const uint32_t primes[7] = {1u, 2654435761u, 805459861u, 3674653429u,
                                       2097192037u, 1434869437u, 2165219737u};
    uint32_t h = 0;
    for (uint32_t d = 0; d < D; ++d) h ^= p[d] * primes[d];
    return h;
}

/* gridencoder.cu:66-84 -- dense stride index while the running stride still fits the level,
 * hash otherwise (gridtype 0 only), then modulo the level size; returns the ENTRY index
 * (the reference multiplies by C and adds the channel; callers here do that). */
static uint32_t orc_grid_index(uint32_t gridtype, int align_corners, uint32_t D, uint32_t hashmap_size,
                               uint32_t resolution, const uint32_t *p) {
    uint32_t stride = 1, index = 0;
    for (uint32_t d = 0; d < D && stride <= hashmap_size; ++d) {
        index += p[d] * stride;
        stride *= align_corners ? resolution : (resolution + 1);
    }
    if (gridtype == 0 && stride > hashmap_size) index = orc_hash(p, D);
    return index % hashmap_size;
}

/* The reference evaluates exp2f(level * S) with the DEVICE math library (<= 2 ulp, not correctly rounded), libm's
 * exp2f can differ in the last bit, which moves fine-level positions by ~1e-4 of a cell.  Tests that compare against a
 * GPU at 1e-5 hand the device's own exp2f values in here (computed with torch.exp2 on the GPU); NULL -> libm. */
static const float *orc_exp2_override = NULL;
static uint32_t orc_exp2_override_n = 0;
void orc_set_level_exp2(const float *values, uint32_t n) { orc_exp2_override = values; orc_exp2_override_n = n; }

/* per-level constants, gridencoder.cu:137-139 (float exp2f, fused `*H - 1`). */
static void orc_level_consts(uint32_t level, float S, uint32_t H, float *scale, uint32_t *resolution) {
    const float e = (orc_exp2_override && level < orc_exp2_override_n) ? orc_exp2_override[level] : exp2f((float)level * S);
    *scale = fmaf(e, (float)H, -1.0f);
    *resolution = (uint32_t)ceilf(*scale) + 1;
}

/* gridencoder.cu:110-117 */
static int orc_oob(const float *x, uint32_t D) {
    for (uint32_t d = 0; d < D; ++d)
        if (x[d] < 0 || x[d] > 1) return 1;
    return 0;
}

/* The reference declares `float pos_deriv[D] = {1.0f}` (gridencoder.cu:143): only element 0 is 1, the
 * rest are 0, so with LINEAR interpolation its dy_dx (and therefore grad_inputs) is zero for d > 0.
 * quirk == 1 (DEFAULT) restates that faithfully -- the oracle is the reference's algorithm, and the B200 kernels
 * reproduce it by default too; quirk == 0 is the mathematically exact derivative, the checker of the kernels'
 * opt-in NGP_OPT_EXACT_INPUT_GRAD / GridEncoder(exact_input_grad=True). */
static int orc_pos_deriv_quirk = 1;
void orc_set_pos_deriv_quirk(int on) { orc_pos_deriv_quirk = on; }

/* gridencoder.cu:146-156: position inside the level; returns base cell + (smooth)fraction. */
static void orc_locate(const float *x, uint32_t D, float scale, int align_corners, uint32_t interp,
                       float *pos, float *pos_deriv, uint32_t *pos_grid) {
    for (uint32_t d = 0; d < D; ++d) {
        float p = fmaf(x[d], scale, align_corners ? 0.0f : 0.5f);
        const float fl = floorf(p);
        pos_grid[d] = (uint32_t)fl;
        p -= (float)pos_grid[d];
        pos_deriv[d] = (orc_pos_deriv_quirk && d > 0) ? 0.0f : 1.0f;
        if (interp == 1) { /* smoothstep, gridencoder.cu:39-47 */
            pos_deriv[d] = 6 * p * (1.0f - p);
            p = p * p * (3.0f - 2.0f * p);
        }
        pos[d] = p;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* hash grid: forward / backward, instantiated for float and double tables                     */
/* ------------------------------------------------------------------------------------------ */
#define ORC_DEFINE_GRID(SUFFIX, T)                                                                       \
/* kernel_grid, gridencoder.cu:87-242.  outputs: [L,B,C] level-major; dy_dx: [B,L,D,C] or NULL */      \
void orc_grid_encode_forward_##SUFFIX(const float *inputs, const T *embeddings, const int *offsets,     \
                                      T *outputs, uint32_t B, uint32_t D, uint32_t C, uint32_t L,       \
                                      float S, uint32_t H, T *dy_dx, uint32_t gridtype,                 \
                                      int align_corners, uint32_t interp) {                             \
    _Pragma("omp parallel for schedule(dynamic, 1)")                                                    \
    for (int64_t level = 0; level < (int64_t)L; ++level) {                                              \
        const T *grid = embeddings + (size_t)(uint32_t)offsets[level] * C;                              \
        const uint32_t hashmap_size = (uint32_t)(offsets[level + 1] - offsets[level]);                  \
        float scale; uint32_t resolution;                                                               \
        orc_level_consts((uint32_t)level, S, H, &scale, &resolution);                                   \
        for (uint32_t b = 0; b < B; ++b) {                                                              \
            const float *x = inputs + (size_t)b * D;                                                    \
            T *out = outputs + ((size_t)level * B + b) * C;                                             \
            T *dd = dy_dx ? dy_dx + ((size_t)b * L + level) * D * C : NULL;                             \
            if (orc_oob(x, D)) {                                                                        \
                for (uint32_t c = 0; c < C; ++c) out[c] = 0;                                            \
                if (dd) for (uint32_t i = 0; i < D * C; ++i) dd[i] = 0;                                 \
                continue;                                                                               \
            }                                                                                           \
            float pos[ORC_MAX_D], pos_deriv[ORC_MAX_D]; uint32_t pg[ORC_MAX_D], pl[ORC_MAX_D];          \
            orc_locate(x, D, scale, align_corners, interp, pos, pos_deriv, pg);                         \
            T res[ORC_MAX_C] = {0};                                                                     \
            for (uint32_t idx = 0; idx < (1u << D); ++idx) {                                            \
                float w = 1;                                                                            \
                for (uint32_t d = 0; d < D; ++d) {                                                      \
                    if ((idx & (1u << d)) == 0) { w *= 1 - pos[d]; pl[d] = pg[d]; }                     \
                    else { w *= pos[d]; pl[d] = pg[d] + 1; }                                            \
                }                                                                                       \
                const size_t e = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,       \
                                                        resolution, pl) * C;                            \
                for (uint32_t c = 0; c < C; ++c) res[c] += (T)w * grid[e + c];                          \
            }                                                                                           \
            for (uint32_t c = 0; c < C; ++c) out[c] = res[c];                                           \
            if (!dd) continue;                                                                          \
            /* analytic d(out)/d(x), gridencoder.cu:198-241 */                                          \
            for (uint32_t gd = 0; gd < D; ++gd) {                                                       \
                T rg[ORC_MAX_C] = {0};                                                                  \
                for (uint32_t idx = 0; idx < (1u << (D - 1)); ++idx) {                                  \
                    float w = scale;                                                                    \
                    for (uint32_t nd = 0; nd < D - 1; ++nd) {                                           \
                        const uint32_t d = (nd >= gd) ? (nd + 1) : nd;                                  \
                        if ((idx & (1u << nd)) == 0) { w *= 1 - pos[d]; pl[d] = pg[d]; }                \
                        else { w *= pos[d]; pl[d] = pg[d] + 1; }                                        \
                    }                                                                                   \
                    pl[gd] = pg[gd];                                                                    \
                    const size_t el = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,  \
                                                             resolution, pl) * C;                       \
                    pl[gd] = pg[gd] + 1;                                                                \
                    const size_t er = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,  \
                                                             resolution, pl) * C;                       \
                    for (uint32_t c = 0; c < C; ++c)                                                    \
                        rg[c] += (T)w * (grid[er + c] - grid[el + c]) * (T)pos_deriv[gd];               \
                }                                                                                       \
                for (uint32_t c = 0; c < C; ++c) dd[gd * C + c] = rg[c];                                \
            }                                                                                           \
        }                                                                                               \
    }                                                                                                   \
}                                                                                                       \
                                                                                                        \
/* kernel_grid_backward + kernel_input_backward, gridencoder.cu:245-366.                            */ \
/* grad: [L,B,C]; grad_embeddings pre-zeroed by the caller (grid.py:89); grad_inputs [B,D] or NULL. */ \
void orc_grid_encode_backward_##SUFFIX(const T *grad, const float *inputs, const T *embeddings,         \
                                       const int *offsets, T *grad_embeddings, uint32_t B, uint32_t D,  \
                                       uint32_t C, uint32_t L, float S, uint32_t H, const T *dy_dx,     \
                                       T *grad_inputs, uint32_t gridtype, int align_corners,            \
                                       uint32_t interp) {                                               \
    (void)embeddings;                                                                                   \
    /* levels own disjoint slices of the table, so level-parallel needs no atomics */                   \
    _Pragma("omp parallel for schedule(dynamic, 1)")                                                    \
    for (int64_t level = 0; level < (int64_t)L; ++level) {                                              \
        T *gg = grad_embeddings + (size_t)(uint32_t)offsets[level] * C;                                 \
        const uint32_t hashmap_size = (uint32_t)(offsets[level + 1] - offsets[level]);                  \
        float scale; uint32_t resolution;                                                               \
        orc_level_consts((uint32_t)level, S, H, &scale, &resolution);                                   \
        for (uint32_t b = 0; b < B; ++b) {                                                              \
            const float *x = inputs + (size_t)b * D;                                                    \
            if (orc_oob(x, D)) continue;                                                                \
            const T *g = grad + ((size_t)level * B + b) * C;                                            \
            float pos[ORC_MAX_D], pos_deriv[ORC_MAX_D]; uint32_t pg[ORC_MAX_D], pl[ORC_MAX_D];          \
            orc_locate(x, D, scale, align_corners, interp, pos, pos_deriv, pg);                         \
            for (uint32_t idx = 0; idx < (1u << D); ++idx) {                                            \
                float w = 1;                                                                            \
                for (uint32_t d = 0; d < D; ++d) {                                                      \
                    if ((idx & (1u << d)) == 0) { w *= 1 - pos[d]; pl[d] = pg[d]; }                     \
                    else { w *= pos[d]; pl[d] = pg[d] + 1; }                                            \
                }                                                                                       \
                const size_t e = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,       \
                                                        resolution, pl) * C;                            \
                for (uint32_t c = 0; c < C; ++c) gg[e + c] += (T)w * g[c];                              \
            }                                                                                           \
        }                                                                                               \
    }                                                                                                   \
    if (dy_dx && grad_inputs) {                                                                         \
        _Pragma("omp parallel for")                                                                     \
        for (int64_t t = 0; t < (int64_t)B * D; ++t) {                                                  \
            const uint32_t b = (uint32_t)(t / D), d = (uint32_t)(t % D);                                \
            const T *dd = dy_dx + (size_t)b * L * D * C;                                                \
            T r = 0;                                                                                    \
            for (uint32_t l = 0; l < L; ++l)                                                            \
                for (uint32_t c = 0; c < C; ++c)                                                        \
                    r += grad[((size_t)l * B + b) * C + c] * dd[(size_t)l * D * C + d * C + c];         \
            grad_inputs[t] = r;                                                                         \
        }                                                                                               \
    }                                                                                                   \
}                                                                                                       \
                                                                                                        \
/* kernel_grad_tv, gridencoder.cu:503-607: adds the normalised TV gradient in place. */                \
void orc_grad_total_variation_##SUFFIX(const T *inputs, const T *embeddings, T *grad, const int *offsets,\
                                       float weight, uint32_t B, uint32_t D, uint32_t C, uint32_t L,    \
                                       float S, uint32_t H, uint32_t gridtype, int align_corners) {     \
    for (uint32_t level = 0; level < L; ++level) {                                                      \
        const T *grid = embeddings + (size_t)(uint32_t)offsets[level] * C;                              \
        T *gg = grad + (size_t)(uint32_t)offsets[level] * C;                                            \
        const uint32_t hashmap_size = (uint32_t)(offsets[level + 1] - offsets[level]);                  \
        float scale; uint32_t resolution;                                                               \
        orc_level_consts(level, S, H, &scale, &resolution);                                             \
        const T w = (T)(weight / (2 * D));                                                              \
        for (uint32_t b = 0; b < B; ++b) {                                                              \
            float x[ORC_MAX_D]; uint32_t pg[ORC_MAX_D];                                                 \
            for (uint32_t d = 0; d < D; ++d) x[d] = (float)inputs[(size_t)b * D + d];                   \
            if (orc_oob(x, D)) continue;                                                                \
            for (uint32_t d = 0; d < D; ++d)                                                            \
                pg[d] = (uint32_t)floorf(fmaf(x[d], scale, align_corners ? 0.0f : 0.5f));               \
            T res[ORC_MAX_C] = {0}, idl[ORC_MAX_C] = {0};                                               \
            const size_t e0 = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,          \
                                                     resolution, pg) * C;                               \
            for (uint32_t d = 0; d < D; ++d) {                                                          \
                const uint32_t cur = pg[d];                                                             \
                if (cur < resolution) {                                                                 \
                    pg[d] = cur + 1;                                                                    \
                    const size_t e = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,   \
                                                            resolution, pg) * C;                        \
                    for (uint32_t c = 0; c < C; ++c) {                                                  \
                        const T v = grid[e0 + c] - grid[e + c]; res[c] += v; idl[c] += v * v; }         \
                }                                                                                       \
                if (cur > 0) {                                                                          \
                    pg[d] = cur - 1;                                                                    \
                    const size_t e = (size_t)orc_grid_index(gridtype, align_corners, D, hashmap_size,   \
                                                            resolution, pg) * C;                        \
                    for (uint32_t c = 0; c < C; ++c) {                                                  \
                        const T v = grid[e0 + c] - grid[e + c]; res[c] += v; idl[c] += v * v; }         \
                }                                                                                       \
                pg[d] = cur;                                                                            \
            }                                                                                           \
            for (uint32_t c = 0; c < C; ++c)                                                            \
                gg[e0 + c] += w * res[c] * (T)(1.0 / sqrt((double)idl[c] + 1e-9));                      \
        }                                                                                               \
    }                                                                                                   \
}

ORC_DEFINE_GRID(f32, float)
ORC_DEFINE_GRID(f64, double)

/* ------------------------------------------------------------------------------------------ */
/* ray marching helpers (raymarching.cu:19-81)                                                 */
/* ------------------------------------------------------------------------------------------ */
#define ORC_SQRT3 1.7320508075688772f
#define ORC_RPI 0.3183098861837907f

static inline float orc_clampf(float x, float lo, float hi) { return fminf(hi, fmaxf(lo, x)); }
static inline float orc_signf(float x) { return copysignf(1.0f, x); }

/* raymarching.cu:56-63: spread the low 10 bits of v so that two zero bits follow each bit */
static inline uint32_t orc_expand_bits(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
static inline uint32_t orc_morton3D_1(uint32_t x, uint32_t y, uint32_t z) {
    return orc_expand_bits(x) | (orc_expand_bits(y) << 1) | (orc_expand_bits(z) << 2);
}
/* raymarching.cu:73-81 */
static inline uint32_t orc_compact_bits(uint32_t x) {
    x = x & 0x49249249;
    x = (x | (x >> 2)) & 0xc30c30c3;
    x = (x | (x >> 4)) & 0x0f00f00f;
    x = (x | (x >> 8)) & 0xff0000ff;
    x = (x | (x >> 16)) & 0x0000ffff;
    return x;
}

/* raymarching.cu:42-47 */
static inline int orc_mip_from_pos(float x, float y, float z, float max_cascade) {
    const float mx = fmaxf(fabsf(x), fmaxf(fabsf(y), fabsf(z)));
    int e;
    frexpf(mx, &e);
    return (int)fminf(max_cascade - 1, fmaxf(0, (float)e));
}
/* raymarching.cu:49-54: `dt * H * 0.5` -- float product, then a DOUBLE multiply, narrowed */
static inline int orc_mip_from_dt(float dt, float H, float max_cascade) {
    const float mx = (float)((double)(dt * H) * 0.5);
    int e;
    frexpf(mx, &e);
    return (int)fminf(max_cascade - 1, fmaxf(0, (float)e));
}

typedef struct {
    float ox, oy, oz, dx, dy, dz, rdx, rdy, rdz;
    float rH, H3, bound, dt_gamma, dt_min, dt_max, Cf, Hf;
    uint32_t H;
    const uint8_t *grid;
} orc_ray_t;

static void orc_ray_init(orc_ray_t *r, const float *o, const float *d, float bound, float dt_gamma,
                         uint32_t max_steps, uint32_t C, uint32_t H, const uint8_t *grid) {
    r->ox = o[0]; r->oy = o[1]; r->oz = o[2];
    r->dx = d[0]; r->dy = d[1]; r->dz = d[2];
    r->rdx = 1 / r->dx; r->rdy = 1 / r->dy; r->rdz = 1 / r->dz;
    r->rH = 1 / (float)H;
    r->H3 = (float)(uint32_t)(H * H * H);
    r->bound = bound; r->dt_gamma = dt_gamma;
    r->dt_min = 2 * ORC_SQRT3 / (float)max_steps;                 /* raymarching.cu:345 */
    r->dt_max = 2 * ORC_SQRT3 * (float)(1 << (C - 1)) / (float)H; /* raymarching.cu:346 */
    r->Cf = (float)C; r->Hf = (float)H; r->H = H; r->grid = grid;
}

/* One iteration of the marching loop body (raymarching.cu:360-399 == :428-478 == :760-812).
 * Returns 1 when the cell at t is occupied (x,y,z,dt valid, t NOT yet advanced), else advances
 * t past the empty voxel and returns 0. */
static inline int orc_march_probe(const orc_ray_t *r, float *t, float *x, float *y, float *z, float *dt) {
    const float tt0 = *t;
    *x = orc_clampf(fmaf(tt0, r->dx, r->ox), -r->bound, r->bound);
    *y = orc_clampf(fmaf(tt0, r->dy, r->oy), -r->bound, r->bound);
    *z = orc_clampf(fmaf(tt0, r->dz, r->oz), -r->bound, r->bound);
    *dt = orc_clampf(tt0 * r->dt_gamma, r->dt_min, r->dt_max);

    const int la = orc_mip_from_pos(*x, *y, *z, r->Cf), lb = orc_mip_from_dt(*dt, r->Hf, r->Cf);
    const int level = la > lb ? la : lb;
    const float mip_bound = fminf(scalbnf(1.0f, level), r->bound);
    const float mip_rbound = 1 / mip_bound;

    /* `0.5 * (x * mip_rbound + 1) * H`: float fma, promoted to double by the literal 0.5,
     * narrowed to float by clamp()'s parameter, truncated to int (raymarching.cu:374-376) */
    const float hi = (float)(r->H - 1);
    const int nx = (int)orc_clampf((float)(0.5 * (double)fmaf(*x, mip_rbound, 1.0f) * (double)r->H), 0.0f, hi);
    const int ny = (int)orc_clampf((float)(0.5 * (double)fmaf(*y, mip_rbound, 1.0f) * (double)r->H), 0.0f, hi);
    const int nz = (int)orc_clampf((float)(0.5 * (double)fmaf(*z, mip_rbound, 1.0f) * (double)r->H), 0.0f, hi);

    /* index arithmetic happens in FLOAT (H3 is a float, raymarching.cu:339,378) */
    const uint32_t index = (uint32_t)fmaf((float)level, r->H3,
                                          (float)orc_morton3D_1((uint32_t)nx, (uint32_t)ny, (uint32_t)nz));
    const int occ = r->grid[index / 8] & (1 << (index % 8));
    if (occ) return 1;

    /* distance to the far face of this voxel, then step until past it (raymarching.cu:390-398) */
    const float tx = fmaf(fmaf(((float)nx + 0.5f + 0.5f * orc_signf(r->dx)) * r->rH, 2.0f, -1.0f), mip_bound, -*x) * r->rdx;
    const float ty = fmaf(fmaf(((float)ny + 0.5f + 0.5f * orc_signf(r->dy)) * r->rH, 2.0f, -1.0f), mip_bound, -*y) * r->rdy;
    const float tz = fmaf(fmaf(((float)nz + 0.5f + 0.5f * orc_signf(r->dz)) * r->rH, 2.0f, -1.0f), mip_bound, -*z) * r->rdz;
    const float tt = tt0 + fmaxf(0.0f, fminf(tx, fminf(ty, tz)));
    float tc = tt0;
    do {
        tc += orc_clampf(tc * r->dt_gamma, r->dt_min, r->dt_max);
    } while (tc < tt);
    *t = tc;
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* utils: raymarching.cu:91-289                                                                */
/* ------------------------------------------------------------------------------------------ */
void orc_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, uint32_t N,
                            float min_near, float *nears, float *fars) {
    for (uint32_t n = 0; n < N; ++n) {
        const float *o = rays_o + 3 * (size_t)n, *d = rays_d + 3 * (size_t)n;
        const float rdx = 1 / d[0], rdy = 1 / d[1], rdz = 1 / d[2];
        float near = (aabb[0] - o[0]) * rdx, far = (aabb[3] - o[0]) * rdx, tmp;
        if (near > far) { tmp = near; near = far; far = tmp; }
        float ny = (aabb[1] - o[1]) * rdy, fy = (aabb[4] - o[1]) * rdy;
        if (ny > fy) { tmp = ny; ny = fy; fy = tmp; }
        if (near > fy || ny > far) { nears[n] = fars[n] = FLT_MAX; continue; }
        if (ny > near) near = ny;
        if (fy < far) far = fy;
        float nz = (aabb[2] - o[2]) * rdz, fz = (aabb[5] - o[2]) * rdz;
        if (nz > fz) { tmp = nz; nz = fz; fz = tmp; }
        if (near > fz || nz > far) { nears[n] = fars[n] = FLT_MAX; continue; }
        if (nz > near) near = nz;
        if (fz < far) far = fz;
        if (near < min_near) near = min_near;
        nears[n] = near; fars[n] = far;
    }
}

/* raymarching.cu:162-198 (tolerance-compared, not bit-exact: sqrt/atan2 library functions) */
void orc_sph_from_ray(const float *rays_o, const float *rays_d, float radius, uint32_t N, float *coords) {
    for (uint32_t n = 0; n < N; ++n) {
        const float *o = rays_o + 3 * (size_t)n, *d = rays_d + 3 * (size_t)n;
        const double A = (double)d[0] * d[0] + (double)d[1] * d[1] + (double)d[2] * d[2];
        const double Bh = (double)o[0] * d[0] + (double)o[1] * d[1] + (double)o[2] * d[2];
        const double Cc = (double)o[0] * o[0] + (double)o[1] * o[1] + (double)o[2] * o[2] - (double)radius * radius;
        const double t = (-Bh + sqrt(Bh * Bh - A * Cc)) / A;
        const double x = o[0] + t * d[0], y = o[1] + t * d[1], z = o[2] + t * d[2];
        coords[2 * (size_t)n + 0] = (float)(2 * atan2(sqrt(x * x + z * z), y) * (double)ORC_RPI - 1);
        coords[2 * (size_t)n + 1] = (float)(atan2(z, x) * (double)ORC_RPI);
    }
}

void orc_morton3D(const int *coords, uint32_t N, int *indices) {
    for (uint32_t n = 0; n < N; ++n)
        indices[n] = (int)orc_morton3D_1((uint32_t)coords[3 * (size_t)n], (uint32_t)coords[3 * (size_t)n + 1],
                                         (uint32_t)coords[3 * (size_t)n + 2]);
}

void orc_morton3D_invert(const int *indices, uint32_t N, int *coords) {
    for (uint32_t n = 0; n < N; ++n) {
        const int ind = indices[n]; /* signed shifts, as the reference (raymarching.cu:249-253) */
        coords[3 * (size_t)n + 0] = (int)orc_compact_bits((uint32_t)(ind >> 0));
        coords[3 * (size_t)n + 1] = (int)orc_compact_bits((uint32_t)(ind >> 1));
        coords[3 * (size_t)n + 2] = (int)orc_compact_bits((uint32_t)(ind >> 2));
    }
}

/* raymarching.cu:267-289: strict `>`; bit i of byte n <- cell 8n+i */
void orc_packbits(const float *grid, uint32_t N, float density_thresh, uint8_t *bitfield) {
    for (uint32_t n = 0; n < N; ++n) {
        uint8_t bits = 0;
        for (int i = 0; i < 8; ++i) bits |= (grid[8 * (size_t)n + i] > density_thresh) ? (uint8_t)(1u << i) : 0;
        bitfield[n] = bits;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* training marcher: raymarching.cu:311-480                                                    */
/* Sample placement in the reference is decided by atomicAdd arrival order (:405-406).  This    */
/* restatement (and the B200 kernel) uses the canonical order: slot == ray id, sample offsets   */
/* == exclusive prefix sum of the per-ray counts.  Per-ray content is identical.                */
/* ------------------------------------------------------------------------------------------ */
void orc_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound,
                          float dt_gamma, uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M,
                          const float *nears, const float *fars, float *xyzs, float *dirs, float *deltas,
                          int *rays, int *counter, const float *noises) {
    uint32_t point_index = (uint32_t)counter[0];
    const uint32_t ray_base = (uint32_t)counter[1];
    for (uint32_t n = 0; n < N; ++n) {
        orc_ray_t r;
        orc_ray_init(&r, rays_o + 3 * (size_t)n, rays_d + 3 * (size_t)n, bound, dt_gamma, max_steps, C, H, grid);
        const float far = fars[n];
        float t0 = nears[n];
        t0 = fmaf(orc_clampf(t0 * dt_gamma, r.dt_min, r.dt_max), noises[n], t0); /* :351 */

        float t = t0, x, y, z, dt;
        uint32_t num_steps = 0;
        while (t < far && num_steps < max_steps) {
            if (orc_march_probe(&r, &t, &x, &y, &z, &dt)) { num_steps++; t += dt; }
        }
        int *rr = rays + 3 * (size_t)n; /* slot == ray id (the reference: atomicAdd(counter+1,1) order) */
        rr[0] = (int)n; rr[1] = (int)point_index; rr[2] = (int)num_steps;
        const uint32_t my_off = point_index;
        point_index += num_steps;
        if (num_steps == 0) continue;
        if (my_off + num_steps > M) continue;

        float *px = xyzs + 3 * (size_t)my_off, *pd = dirs + 3 * (size_t)my_off, *pl = deltas + 2 * (size_t)my_off;
        t = t0;
        float last_t = t;
        uint32_t step = 0;
        while (t < far && step < num_steps) {
            if (orc_march_probe(&r, &t, &x, &y, &z, &dt)) {
                px[0] = x; px[1] = y; px[2] = z;
                pd[0] = r.dx; pd[1] = r.dy; pd[2] = r.dz;
                t += dt;
                pl[0] = dt; pl[1] = t - last_t;
                last_t = t;
                px += 3; pd += 3; pl += 2; step++;
            }
        }
    }
    counter[0] = (int)point_index;
    counter[1] = (int)(ray_base + N);
}

/* raymarching.cu:500-577 */
void orc_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas, const int *rays,
                                      uint32_t M, uint32_t N, float T_thresh, float *weights_sum, float *depth,
                                      float *image) {
    for (uint32_t n = 0; n < N; ++n) {
        const uint32_t index = (uint32_t)rays[3 * (size_t)n], offset = (uint32_t)rays[3 * (size_t)n + 1],
                       num_steps = (uint32_t)rays[3 * (size_t)n + 2];
        float r = 0, g = 0, b = 0, ws = 0, t = 0, d = 0, T = 1.0f;
        if (!(num_steps == 0 || offset + num_steps > M)) {
            for (uint32_t s = 0; s < num_steps; ++s) {
                const size_t i = (size_t)offset + s;
                const float alpha = 1.0f - expf(-sigmas[i] * deltas[2 * i]);
                const float w = alpha * T;
                r += w * rgbs[3 * i]; g += w * rgbs[3 * i + 1]; b += w * rgbs[3 * i + 2];
                t += deltas[2 * i + 1];
                d += w * t;
                ws += w;
                T *= 1.0f - alpha;
                if (T < T_thresh) break;
            }
        }
        weights_sum[index] = ws; depth[index] = d;
        image[3 * (size_t)index] = r; image[3 * (size_t)index + 1] = g; image[3 * (size_t)index + 2] = b;
    }
}

/* raymarching.cu:601-691; grad_sigmas / grad_rgbs pre-zeroed by the caller (raymarching.py:284-285) */
void orc_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_depth, const float *grad_image,
                                       const float *sigmas, const float *rgbs, const float *deltas, const int *rays,
                                       const float *weights_sum, const float *depth, const float *image, uint32_t M,
                                       uint32_t N, float T_thresh, float *grad_sigmas, float *grad_rgbs) {
    for (uint32_t n = 0; n < N; ++n) {
        const uint32_t index = (uint32_t)rays[3 * (size_t)n], offset = (uint32_t)rays[3 * (size_t)n + 1],
                       num_steps = (uint32_t)rays[3 * (size_t)n + 2];
        if (num_steps == 0 || offset + num_steps > M) continue;
        const float *gi = grad_image + 3 * (size_t)index;
        const float gws = grad_weights_sum[index], gd = grad_depth[index];
        const float rf = image[3 * (size_t)index], gf = image[3 * (size_t)index + 1], bf = image[3 * (size_t)index + 2];
        const float wsf = weights_sum[index], df = depth[index];
        float r = 0, g = 0, b = 0, ws = 0, t = 0, d = 0, T = 1.0f;
        for (uint32_t s = 0; s < num_steps; ++s) {
            const size_t i = (size_t)offset + s;
            const float alpha = 1.0f - expf(-sigmas[i] * deltas[2 * i]);
            const float w = alpha * T;
            r += w * rgbs[3 * i]; g += w * rgbs[3 * i + 1]; b += w * rgbs[3 * i + 2];
            t += deltas[2 * i + 1];
            d += w * t;
            ws += w;
            T *= 1.0f - alpha;
            grad_rgbs[3 * i] = gi[0] * w; grad_rgbs[3 * i + 1] = gi[1] * w; grad_rgbs[3 * i + 2] = gi[2] * w;
            grad_sigmas[i] = deltas[2 * i] * (gi[0] * (T * rgbs[3 * i] - (rf - r)) + gi[1] * (T * rgbs[3 * i + 1] - (gf - g)) +
                                              gi[2] * (T * rgbs[3 * i + 2] - (bf - b)) + gd * (T * t - (df - d)) +
                                              gws * (1 - wsf));
            if (T < T_thresh) break;
        }
    }
}

/* ------------------------------------------------------------------------------------------ */
/* inference pair: raymarching.cu:709-914                                                      */
/* ------------------------------------------------------------------------------------------ */
void orc_march_rays(uint32_t n_alive, uint32_t n_step, const int *rays_alive, const float *rays_t, const float *rays_o,
                    const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t C, uint32_t H,
                    const uint8_t *grid, const float *nears, const float *fars, float *xyzs, float *dirs, float *deltas,
                    const float *noises) {
    (void)nears;
    for (uint32_t n = 0; n < n_alive; ++n) {
        const int index = rays_alive[n];
        orc_ray_t r;
        orc_ray_init(&r, rays_o + 3 * (size_t)index, rays_d + 3 * (size_t)index, bound, dt_gamma, max_steps, C, H, grid);
        float *px = xyzs + 3 * (size_t)n * n_step, *pd = dirs + 3 * (size_t)n * n_step, *pl = deltas + 2 * (size_t)n * n_step;
        float t = rays_t[index];
        const float far = fars[index];
        t = fmaf(orc_clampf(t * dt_gamma, r.dt_min, r.dt_max), noises[n], t); /* :755 */
        float last_t = t, x, y, z, dt;
        uint32_t step = 0;
        while (t < far && step < n_step) {
            if (orc_march_probe(&r, &t, &x, &y, &z, &dt)) {
                px[0] = x; px[1] = y; px[2] = z;
                pd[0] = r.dx; pd[1] = r.dy; pd[2] = r.dz;
                t += dt;
                pl[0] = dt; pl[1] = t - last_t;
                last_t = t;
                px += 3; pd += 3; pl += 2; step++;
            }
        }
    }
}

void orc_composite_rays(uint32_t n_alive, uint32_t n_step, float T_thresh, int *rays_alive, float *rays_t,
                        const float *sigmas, const float *rgbs, const float *deltas, float *weights_sum, float *depth,
                        float *image) {
    for (uint32_t n = 0; n < n_alive; ++n) {
        const int index = rays_alive[n];
        float t = rays_t[index], ws = weights_sum[index], d = depth[index];
        float r = image[3 * (size_t)index], g = image[3 * (size_t)index + 1], b = image[3 * (size_t)index + 2];
        uint32_t step = 0;
        while (step < n_step) {
            const size_t i = (size_t)n * n_step + step;
            if (deltas[2 * i] == 0) break;
            const float alpha = 1.0f - expf(-sigmas[i] * deltas[2 * i]);
            const float T = 1 - ws; /* :877 */
            const float w = alpha * T;
            ws += w;
            t += deltas[2 * i + 1];
            d += w * t;
            r += w * rgbs[3 * i]; g += w * rgbs[3 * i + 1]; b += w * rgbs[3 * i + 2];
            if (T < T_thresh) break;
            step++;
        }
        if (step < n_step) rays_alive[n] = -1; else rays_t[index] = t;
        weights_sum[index] = ws; depth[index] = d;
        image[3 * (size_t)index] = r; image[3 * (size_t)index + 1] = g; image[3 * (size_t)index + 2] = b;
    }
}
