/*
 * ngp_b200.h -- C ABI of libngp_b200.so, the B200 (sm_100a) implementation of torch-ngp's
 * Instant-NGP hot path.  This is the drop-in boundary: one `extern "C"` entry point for each
 * free function that the reference's three pybind11 modules export
 *   gridencoder : /root/reference/gridencoder/src/gridencoder.h:12-15 , bindings.cpp:5-9
 *   raymarching : /root/reference/raymarching/src/raymarching.h:7-18 , bindings.cpp:5-18
 *   ffmlp       : /root/reference/ffmlp/src/ffmlp.h:8-14             , bindings.cpp:5-10
 * with `at::Tensor` arguments replaced by raw DEVICE pointers (caller-allocated, caller-owned,
 * contiguous, exactly the shapes/dtypes the reference documents) and the CUDA stream made an
 * explicit last argument (the reference launches on the legacy default stream).
 *
 * Conventions
 *   - every function returns 0 on success, a negative NGP_E* code otherwise; the message for the
 *     calling thread's last failure is ngp_last_error().
 *   - no allocation, no host synchronisation and no hidden streams inside the library: all work
 *     is enqueued on `stream` and is CUDA-graph capturable.
 *   - `dtype` selects the table / activation element type where the reference dispatches with
 *     AT_DISPATCH_FLOATING_TYPES_AND_HALF.
 *   - `void *stream` is a cudaStream_t (0 == legacy default stream).
 */
#ifndef NGP_B200_H
#define NGP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define NGP_API __attribute__((visibility("default")))
#else
#define NGP_API
#endif

#define NGP_OK 0
#define NGP_EINVAL (-1)      /* bad argument (unsupported D/C/hidden_dim, null pointer, ...) */
#define NGP_ECUDA (-2)       /* a CUDA runtime call / launch failed                           */
#define NGP_EUNSUPPORTED (-3)

#define NGP_DTYPE_F32 0
#define NGP_DTYPE_F16 1
#define NGP_DTYPE_F64 2

/* Memory layout of the encoder's per-point feature matrix (outputs / grad):
 *   NGP_LAYOUT_LBC : [L, B, C] level-major -- what the reference kernels use (gridencoder.cu:385)
 *   NGP_LAYOUT_BLC : [B, L*C]  point-major -- what GridEncoder.forward returns (grid.py:69); using it
 *                    removes the two transposing copies of grid.py:69 and grid.py:87.            */
#define NGP_LAYOUT_LBC 0
#define NGP_LAYOUT_BLC 1

NGP_API const char *ngp_last_error(void);
/* 2 since ngp_call_opts carries level_begin / level_end (a caller built against version 1 passes a shorter struct: check this first) */
NGP_API int ngp_abi_version(void);
/* Development knobs used by bench/tests to A/B kernel VARIANTS that compute the same results (e.g. "hashgrid_fwd_variant",
 * "march_variant") and to switch the per-launch event profile on; unknown keys -> NGP_EINVAL.  Nothing that changes
 * what a call computes or which rows it touches lives here: that is ngp_call_opts, passed with each call. */
NGP_API int ngp_set_option(const char *key, int value);
NGP_API int ngp_get_option(const char *key, int *value);
/* Launch accounting: total kernel launches made by the library in this process, and -- while option
 * "profile_events" is 1 (not during CUDA-graph capture) -- per-kernel CUDA-event durations, returned as
 * text lines "<kernel> <count> <total_ms>\n" (the call synchronises on the recorded events and clears them). */
NGP_API long long ngp_launch_count(void);
/* Per-call options of the per-sample entry points (the trailing `const ngp_call_opts *opts` argument; NULL = defaults).
 * There is NO process-wide state behind these: every call carries its own, so modules, threads and streams never see
 * each other's settings (two trainers and a bare GridEncoder may run concurrently).  The struct is read on the host
 * while the call enqueues its kernels; it may live on the caller's stack.
 *   batch_limit : device int32 (or NULL).  The kernels of the call process only the first min(B, *batch_limit) rows;
 *                 *batch_limit is read ON THE DEVICE when the kernel runs, so the sample counter that
 *                 ngp_march_rays_train just wrote (counter[0]) bounds the rest of a captured step without a host
 *                 read-back.  Rows past the limit are neither read nor written; strides keep using B.  The reference
 *                 sizes these launches from a host copy of the counter (raymarching.py:221-226, a D2H sync).
 *   input_bound : > 0: `inputs` of ngp_grid_encode_forward / _backward are positions in [-bound, bound]^3 and the
 *                 kernels apply the module's own mapping (x + bound) / (2 bound) (gridencoder/grid.py:807) as they
 *                 load them (D = 3, C = 2 fast kernels only: NGP_EUNSUPPORTED otherwise).  <= 0: inputs in [0, 1].
 *   flags       : NGP_OPT_* below.
 *   grid_cap_blocks_per_sm : > 0 caps the grid of ngp_adam_step at that many blocks per SM (grid-stride), so that the
 *                 HBM-bound optimizer can share the SMs with another stream's kernels; 0 = one block per 1024 items.
 *   level_begin, level_end : ngp_grid_encode_backward only.  level_end > 0: accumulate the table gradient of levels
 *                 [level_begin, level_end) alone (multiples of 4, D = 3, C = 2 fast kernels: NGP_EUNSUPPORTED otherwise;
 *                 grad_inputs must be NULL).  A caller that all-reduces the table gradient across GPUs issues the
 *                 backward as a few level ranges, finest first, and starts the collective of each range while the next
 *                 one is still being accumulated (nerf/sdf_trainer.py).  Both 0 = all levels in one call. */
typedef struct ngp_call_opts {
    const int32_t *batch_limit;
    float input_bound;
    uint32_t flags;
    int32_t grid_cap_blocks_per_sm;
    uint32_t level_begin, level_end;
} ngp_call_opts;
/* dy_dx of ngp_grid_encode_forward with linear interpolation: the reference initialises `float pos_deriv[D] = {1.0f}`
 * (gridencoder.cu:143), i.e. only d = 0 gets a non-zero input derivative.  DEFAULT = the reference's values, bit for
 * bit.  This flag selects the mathematically exact derivative (all D components) instead. */
#define NGP_OPT_EXACT_INPUT_GRAD 1u
/* ngp_grid_encode_backward: accumulate the levels that fit in shared memory there (fp32, one flush per block) instead
 * of reducing them in L2 -- for UNORDERED points (SDF batches, random points), where every point lands in one of a few
 * thousand coarse cells.  Same sums up to addition order. */
#define NGP_OPT_GRID_BWD_SMEM_COARSE 2u
/* ngp_grid_encode_backward, fp16 table: always ONE 16-byte reduction (red.v4.f16x2) per aligned 4-entry group that holds both
 * x-neighbours of a corner row, also for the lanes that carry a merged run of points.  Default: decided per warp (ray-ordered
 * samples are faster without: +20 % with it).  For batches ordered by ngp_bin_points it is 11 % faster (452 vs 510 us per
 * 2^20 torus samples, same job).  Same sums. */
#define NGP_OPT_GRID_BWD_GROUP_RED 4u
NGP_API int ngp_profile_collect(char *buf, size_t buf_bytes);

/* ---------------------------------------------------------------------------------------------
 * gridencoder  (replaces gridencoder.h:12-15)
 * ------------------------------------------------------------------------------------------- */
/* grid_encode_forward (gridencoder.cu:445-468).  inputs [B,D] f32 in [0,1]; embeddings [T,C] dtype;
 * offsets [L+1] i32 (device); offsets_host: the same L+1 values in HOST memory, or NULL -- lets the
 * library plan the shared-memory-resident coarse levels and wide accesses without a device read
 * (NULL selects the generic kernels); outputs dtype in `layout`; dy_dx [B,L,D,C] dtype or NULL.
 * D in {2,3,4,5}, C in {1,2,4,8} ("GridEncoding: C must be 1, 2, 4, or 8." otherwise). */
NGP_API int ngp_grid_encode_forward(const float *inputs, const void *embeddings, const int *offsets, const int *offsets_host,
                            void *outputs, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H, void *dy_dx,
                            uint32_t gridtype, int align_corners, uint32_t interp, int dtype, int layout, void *stream, const ngp_call_opts *opts);
/* grid_encode_backward (gridencoder.cu:470-500).  grad dtype in `layout`; grad_embeddings [T,C] dtype,
 * ACCUMULATED into (caller zero-fills, grid.py:89); dy_dx/grad_inputs ([B,D] dtype) both NULL or both set. */
NGP_API int ngp_grid_encode_backward(const void *grad, const float *inputs, const void *embeddings, const int *offsets,
                             const int *offsets_host, void *grad_embeddings, uint32_t B, uint32_t D, uint32_t C, uint32_t L,
                             float S, uint32_t H, const void *dy_dx, void *grad_inputs, uint32_t gridtype, int align_corners,
                             uint32_t interp, int dtype, int layout, void *stream, const ngp_call_opts *opts);
/* grad_total_variation (gridencoder.cu:636-642).  inputs [B,D] in the TABLE dtype (sic, :505). */
NGP_API int ngp_grad_total_variation(const void *inputs, const void *embeddings, void *grad, const int *offsets,
                             float weight, uint32_t B, uint32_t D, uint32_t C, uint32_t L, float S, uint32_t H,
                             uint32_t gridtype, int align_corners, int dtype, void *stream);

/* ---------------------------------------------------------------------------------------------
 * raymarching  (replaces raymarching.h:7-18).  All floating tensors are f32 (the Python layer
 * forces it: raymarching.py:21,54,131,163,240,300,354).
 * ------------------------------------------------------------------------------------------- */
NGP_API int ngp_near_far_from_aabb(const float *rays_o, const float *rays_d, const float *aabb, uint32_t N, float min_near,
                           float *nears, float *fars, void *stream);
NGP_API int ngp_sph_from_ray(const float *rays_o, const float *rays_d, float radius, uint32_t N, float *coords, void *stream);
NGP_API int ngp_morton3D(const int *coords, uint32_t N, int *indices, void *stream);
NGP_API int ngp_morton3D_invert(const int *indices, uint32_t N, int *coords, void *stream);
/* NeRFRenderer.mark_untrained_grid (nerf/renderer.py:433-496; the reference's 5-deep Python loop over cell blocks,
 * cascades and camera batches) as one kernel.  poses [B,4,4] f32 camera-to-world (device), intrinsics fx fy cx cy;
 * density_grid [C, H^3] in Morton order: cells whose centre no camera sees (frustum widened by one cell) become -1.
 * count_out [C, H^3] i32 or NULL: the per-cell camera count of the reference's `count` tensor. */
NGP_API int ngp_mark_untrained_grid(const float *poses, uint32_t B, double fx, double fy, double cx, double cy, float bound, uint32_t C,
                            uint32_t H, float *density_grid, int *count_out, void *stream);
/* N = number of output BYTES = C*H^3/8 (raymarching.cu:292) */
NGP_API int ngp_packbits(const float *grid, uint32_t N, float density_thresh, uint8_t *bitfield, void *stream);

/* march_rays_train (raymarching.cu:482-490).  Differences from the reference, by design:
 *   - sample placement is canonical instead of atomicAdd-ordered (:405-406): rays[n] = (n, offset_n, count_n)
 *     with offset_n = counter[0] + exclusive prefix sum of the counts in ray order (per-ray content is
 *     bit-identical; counter[0] += total, counter[1] += N exactly as the reference leaves them);
 *   - needs `workspace` of ngp_march_rays_train_workspace_bytes(N) bytes (chained-scan state). */
NGP_API size_t ngp_march_rays_train_workspace_bytes(uint32_t N);
/* larger workspace that also holds per-ray emit masks: pass 2 of the marcher replays them instead of re-testing the
 * occupancy grid (optional; with the smaller workspace above the samples are re-marched) */
NGP_API size_t ngp_march_rays_train_workspace_bytes_ex(uint32_t N, float bound, uint32_t max_steps);
NGP_API int ngp_march_rays_train(const float *rays_o, const float *rays_d, const uint8_t *grid, float bound, float dt_gamma,
                         uint32_t max_steps, uint32_t N, uint32_t C, uint32_t H, uint32_t M, const float *nears,
                         const float *fars, float *xyzs, float *dirs, float *deltas, int *rays, int *counter,
                         const float *noises, void *workspace, size_t workspace_bytes, void *stream);
NGP_API int ngp_composite_rays_train_forward(const float *sigmas, const float *rgbs, const float *deltas, const int *rays,
                                     uint32_t M, uint32_t N, float T_thresh, float *weights_sum, float *depth,
                                     float *image, void *stream);
NGP_API int ngp_composite_rays_train_backward(const float *grad_weights_sum, const float *grad_depth, const float *grad_image,
                                      const float *sigmas, const float *rgbs, const float *deltas, const int *rays,
                                      const float *weights_sum, const float *depth, const float *image, uint32_t M,
                                      uint32_t N, float T_thresh, float *grad_sigmas, float *grad_rgbs, void *stream);
/* march_rays / composite_rays (raymarching.cu:817-824, :917-923).  n_alive_dev (nullable): device int that
 * further clamps n_alive inside the kernel, so a render loop can run without reading the alive count back. */
NGP_API int ngp_march_rays(uint32_t n_alive, const int *n_alive_dev, uint32_t n_step, const int *rays_alive, const float *rays_t,
                   const float *rays_o, const float *rays_d, float bound, float dt_gamma, uint32_t max_steps, uint32_t C,
                   uint32_t H, const uint8_t *grid, const float *nears, const float *fars, float *xyzs, float *dirs,
                   float *deltas, const float *noises, void *stream);
NGP_API int ngp_composite_rays(uint32_t n_alive, const int *n_alive_dev, uint32_t n_step, float T_thresh, int *rays_alive,
                       float *rays_t, const float *sigmas, const float *rgbs, const float *deltas, float *weights_sum,
                       float *depth, float *image, void *stream);
/* Device-side replacement of `rays_alive = rays_alive[rays_alive >= 0]` (nerf/renderer.py:416): stable
 * compaction of the non-negative ids of in[0..min(n,*n_dev)) into `out` (!= in); count -> *n_out (device
 * int, may alias n_dev). */
NGP_API size_t ngp_compact_alive_workspace_bytes(uint32_t n);
NGP_API int ngp_compact_alive(const int *in, uint32_t n, const int *n_dev, int *out, int *n_out, void *workspace,
                      size_t workspace_bytes, void *stream);
/* The same compaction for callers that cannot allocate (code under CUDA-graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  NO counterpart in the reference.  Takes no
 * workspace: the scan scratch is a file-scope `static __device__` array inside the library (about 16 KB, held by every
 * GPU that loads the library until the process exits), so the call performs no cudaMalloc / cudaFree / allocating
 * library routine.  n > NGP_COMPACT_STATIC_MAX_N -> NGP_EINVAL with a message naming the limit.  Results are identical
 * to ngp_compact_alive.  All calls on one device share that array: calls that overlap on the GPU (different streams or
 * threads) race -- serialise them on one stream or use ngp_compact_alive with a workspace per stream.  in != out;
 * n_out may alias n_dev. */
#define NGP_COMPACT_STATIC_MAX_N 4194304 /* 2^22 ids: two 1920x1080 frames' worth of rays */
NGP_API int ngp_compact_alive_static(const int *in, uint32_t n, const int *n_dev, int *out, int *n_out, void *stream);

/* ---------------------------------------------------------------------------------------------
 * ffmlp  (replaces ffmlp.h:8-14).  fp16 activations, fp32 accumulation on tcgen05 tensor cores (TMEM).
 * Row-major [B, dim] activations (== the reference's "col-major [dim, B]"), weights flat, each layer
 * row-major [out, in] (ffmlp.cu:631-634), fp16 or fp32 (`weights_dtype`; fp32 is rounded to fp16 while
 * staged into shared memory).  Any B >= 0 (the ragged tail tile is masked; FFMLP.forward's padding copy,
 * ffmlp.py:157-159, is not needed).  hidden_dim in {16,32,64,128,256} (ffmlp.cu:651-658), input_dim % 16 == 0
 * (<= 128; <= 256 at hidden_dim 256), output_dim % 16 == 0 (<= 128), num_layers >= 2 (num_layers + 1 matmuls), activation = any of the reference's
 * codes 0..6 (ffmlp.cu:22-32: ReLU, Exponential, Sine, Sigmoid, Squareplus, Softplus, None; Sine has no backward, as
 * in the reference, utils.h:552-556), output_activation 6 = None (all that FFMLP ever passes, ffmlp.py:107-108);
 * anything else -> NGP_EUNSUPPORTED.
 * hidden_dim <= 128: forward_buffer / inference_buffer / backward_buffer are accepted for signature compatibility
 * and may be NULL: they are never written (activations live in shared memory; the backward recomputes the forward).
 * hidden_dim 256 (neither the weights nor the weight-gradient accumulators fit on the SM next to a batch tile): the
 * reference's own two-phase flow (ffmlp.cu:749-895) and buffers -- ngp_ffmlp_forward writes forward_buffer
 * [num_layers, B, 256] when it is not NULL, ngp_ffmlp_backward REQUIRES that forward_buffer and a backward_buffer of
 * the same shape (scratch; any contents), NGP_EINVAL otherwise; inference_buffer stays unused.
 * ------------------------------------------------------------------------------------------- */
NGP_API int ngp_ffmlp_forward(const void *inputs, const void *weights, int weights_dtype, uint32_t B, uint32_t input_dim,
                      uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                      uint32_t output_activation, void *forward_buffer, void *outputs, void *stream, const ngp_call_opts *opts);
NGP_API int ngp_ffmlp_inference(const void *inputs, const void *weights, int weights_dtype, uint32_t B, uint32_t input_dim,
                        uint32_t output_dim, uint32_t hidden_dim, uint32_t num_layers, uint32_t activation,
                        uint32_t output_activation, void *inference_buffer, void *outputs, void *stream, const ngp_call_opts *opts);
/* grad_weights: fp16 like the reference (NGP_DTYPE_F16) or fp32 (NGP_DTYPE_F32); OVERWRITTEN, not accumulated.
 * grad_inputs [B, input_dim] fp16, written iff calc_grad_inputs.  workspace: ngp_ffmlp_backward_workspace_bytes()
 * bytes, zero-initialised once by the caller; the kernel leaves it zeroed for the next call. */
NGP_API int ngp_ffmlp_backward(const void *grad, const void *inputs, const void *weights, int weights_dtype,
                       const void *forward_buffer, uint32_t B, uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim,
                       uint32_t num_layers, uint32_t activation, uint32_t output_activation, int calc_grad_inputs,
                       void *backward_buffer, void *grad_inputs, void *grad_weights, int grad_weights_dtype, void *workspace,
                       size_t workspace_bytes, void *stream, const ngp_call_opts *opts);
NGP_API size_t ngp_ffmlp_backward_workspace_bytes(uint32_t input_dim, uint32_t output_dim, uint32_t hidden_dim,
                                          uint32_t num_layers);
/* The reference's process-global split-k stream pool (ffmlp.cu:711-740) has no equivalent: no-ops. */
NGP_API int ngp_allocate_splitk(size_t size);
NGP_API int ngp_free_splitk(void);

/* ---------------------------------------------------------------------------------------------
 * "next" rows of the scope table (callers either side of the path)
 * ------------------------------------------------------------------------------------------- */
/* SH direction encoding, /root/reference/shencoder/src/shencoder.cu:27-383 (degree 1..4 here).
 * inputs [B,3] f32; outputs [B, out_stride] dtype (first degree^2 columns written; out_stride >= degree^2
 * lets the caller encode straight into a wider, padded MLP input). */
NGP_API int ngp_sh_encode_forward(const float *inputs, void *outputs, uint32_t B, uint32_t degree, uint32_t out_stride,
                          int dtype, void *stream);
/* sh_encode_backward (shencoder.cu:366-383 contracted with the dy_dx of :122-250 in one pass, no dy_dx buffer):
 * grad [B, grad_stride] f32 (first degree^2 columns used), inputs [B,3] f32 -> grad_inputs [B,3] f32. */
NGP_API int ngp_sh_encode_backward(const float *grad, const float *inputs, float *grad_inputs, uint32_t B, uint32_t degree,
                           uint32_t grad_stride, void *stream);

/* Renderer-driver glue (scope table "next" row 1: nerf/renderer.py:257-431 run_cuda, :498-598 update_extra_state;
 * network nerf/network_ff.py:51-74; optimizer step main_nerf.py:194-197 + nerf/utils.py:974-978).  All device
 * resident, no host reads; `state` is a float[16] device block (see nerf_step.cu) holding step count, lr, bias
 * corrections, loss scale, found_inf flag and the loss accumulator. */
NGP_API int ngp_nerf_x01(const float *xyzs, float bound, float *x01, uint32_t M, void *stream, const ngp_call_opts *opts);
/* h [M,16] f16, dirs [M,3] -> sigmas [M] f32 = density_scale*exp(h0); cin [M,32] f16 = [SH4(dir) | h1..h15 | 0] */
NGP_API int ngp_nerf_mid_forward(const void *h, const float *dirs, float density_scale, float *sigmas, void *cin, uint32_t M, void *stream, const ngp_call_opts *opts);
/* c [M,16] f16 -> rgbs [M,3] f32 = sigmoid(c0..c2) */
NGP_API int ngp_nerf_rgb_forward(const void *c, float *rgbs, uint32_t M, void *stream, const ngp_call_opts *opts);
/* MSE loss of image + (1 - weights_sum) * bg against gt [N,3]; loss accumulated into *loss_out; gradients scaled by *loss_scale */
NGP_API int ngp_nerf_loss(const float *image, const float *weights_sum, const float *gt, float bg, uint32_t N, const float *loss_scale,
                  float *grad_image, float *grad_weights_sum, float *grad_depth, float *loss_out, void *stream);
NGP_API int ngp_nerf_rgb_backward(const float *grad_rgbs, const float *rgbs, void *grad_c, uint32_t M, void *stream, const ngp_call_opts *opts);
NGP_API int ngp_nerf_mid_backward(const float *grad_sigmas, const void *h, const void *grad_cin, float density_scale, void *grad_h,
                          uint32_t M, void *stream, const ngp_call_opts *opts);
/* The two nets of NeRFNetwork(--ff) with the glue between them fused into the tcgen05 MLP kernels (network_ff.py:51-74,
 * activation.py:5-17, shencoder.cu:27-120; csrc/ffmlp.cu "IO" modes): same arithmetic as ngp_ffmlp_* + ngp_nerf_mid_* +
 * ngp_nerf_rgb_* composed, bit for bit, without cin / c / grad_c / grad_cin in HBM and without the four elementwise launches.
 * Topology fixed by the reference: 32 -> 64 x num_layers -> 16, ReLU, fp32 (or fp16) parameters in FFMLP's layout.
 *   sigma_forward : enc [M,32] f16 -> h [M,16] f16 and sigmas [M] f32 = density_scale * exp(h[:,0])
 *   color_forward : dirs [M,3] f32, h [M,16] f16 -> rgbs [M,3] f32 = sigmoid(colour net([SH4(dir) | h[:,1:16] | 0])[:, :3])
 *   color_backward: grad_rgbs, rgbs [M,3], grad_sigmas [M] (from composite_rays_train_backward), dirs, h -> grad_weights of the
 *                   colour net (overwritten; workspace as ngp_ffmlp_backward) and grad_h [M,16] f16 = dL/dh of the sigma net's
 *                   output: column 0 from the density (trunc_exp backward), columns 1..15 through the colour net */
NGP_API int ngp_nerf_sigma_forward(const void *enc, const void *weights, int weights_dtype, uint32_t M, uint32_t num_layers, float density_scale,
                           void *h, float *sigmas, void *stream, const ngp_call_opts *opts);
NGP_API int ngp_nerf_color_forward(const float *dirs, const void *h, const void *weights, int weights_dtype, uint32_t M, uint32_t num_layers,
                           float *rgbs, void *stream, const ngp_call_opts *opts);
/* sigma_forward + color_forward in ONE launch (both weight sets staged once per CTA, the encoding tile becomes the colour net's
 * input tile in place): h may be NULL (inference: the sigma net's output never leaves the SM); non-NULL h [M,16] is written for
 * ngp_nerf_color_backward.  Same results as the two calls, bit for bit. */
NGP_API int ngp_nerf_nets_forward(const void *enc, const float *dirs, const void *weights_sigma, const void *weights_color, int weights_dtype,
                          uint32_t M, uint32_t num_layers_sigma, uint32_t num_layers_color, float density_scale, void *h, float *sigmas,
                          float *rgbs, void *stream, const ngp_call_opts *opts);
NGP_API int ngp_nerf_color_backward(const float *grad_rgbs, const float *rgbs, const float *grad_sigmas, const float *dirs, const void *h,
                            const void *weights, int weights_dtype, uint32_t M, uint32_t num_layers, float density_scale, void *grad_h,
                            void *grad_weights, int grad_weights_dtype, void *workspace, size_t workspace_bytes, void *stream,
                            const ngp_call_opts *opts);
/* SDF regression loss (sdf/utils.py:217-225, loss.py:7-15 mape_loss, mean over B): h [B,16] f16 (column 0 = prediction),
 * target [B] f32 -> *loss_out += loss, grad_h [B,16] f16 (scaled by *loss_scale) */
NGP_API int ngp_sdf_loss(const void *h, const float *target, uint32_t B, const float *loss_scale, void *grad_h, float *loss_out, void *stream, const ngp_call_opts *opts);
/* The SDF net's forward (sdf/netowrk_ff.py:43-50: FFMLP 32 -> 64 x num_layers -> 1, padded to 16) and ngp_sdf_loss in ONE launch for the
 * training step: the prediction never leaves the SM, the kernel writes grad_h [B,16] f16 (bit-identical to ngp_ffmlp_forward +
 * ngp_sdf_loss) and accumulates the loss into *loss_out.  enc [B,32] f16, fp32 parameters in FFMLP's layout. */
NGP_API int ngp_sdf_forward_loss(const void *enc, const void *weights, int weights_dtype, const float *target, uint32_t B, uint32_t num_layers,
                         const float *loss_scale, void *grad_h, float *loss_out, void *stream, const ngp_call_opts *opts);
/* Spatial binning of an unordered batch (no counterpart in the reference, whose SDF batches stay in the random order of
 * sdf/provider.py:60-90): points [B,3] f32 in [-bound, bound]^3 and an optional per-point payload values [B] f32 are
 * COPIED to points_out / values_out in the order of the Morton code of a (2^bits)^3 cube grid, bits in {5, 6, 7}; points
 * of one cube stay in arbitrary order.  The permutation is not returned: the caller's loss is a mean over the batch.
 * workspace: ngp_bin_points_workspace(B, bits) bytes, 16-byte aligned, ZERO-FILLED before the first call (every call
 * leaves its histogram zeroed again).  Four launches (histogram with slot ranks, block sums, exclusive scan, scatter). */
NGP_API int ngp_bin_points(const float *points, const float *values, uint32_t B, float bound, uint32_t bits, void *workspace, size_t workspace_bytes,
                   float *points_out, float *values_out, void *stream);
NGP_API size_t ngp_bin_points_workspace(uint32_t B, uint32_t bits);
NGP_API int ngp_step_begin(float *state, float beta1, float beta2, void *stream);
NGP_API int ngp_step_end(float *state, void *stream);
/* sets state.found_inf when any gradient is inf/nan: g16 = n16 halves (n16 % 8 == 0), g32 = n32 floats (n32 % 4 == 0) */
NGP_API int ngp_found_inf(const void *g16, size_t n16, const float *g32, size_t n32, float *state, void *stream);
/* Adam over params[0 .. n16+n32): first n16 have fp16 gradients + an fp16 shadow copy, last n32 fp32 gradients; gradients are
 * divided by (loss scale * grad_div), consumed and CLEARED; skipped (but still cleared) when state.found_inf;
 * weight_decay32: torch.optim.Adam's L2 weight decay for the fp32-gradient group (main_sdf.py:51-54), 0 for NeRF */
NGP_API int ngp_adam_step(float *params, float *exp_avg, float *exp_avg_sq, void *g16, float *g32, void *shadow16, size_t n16, size_t n32,
                  const float *state, float beta1, float beta2, float eps, float grad_div, float weight_decay32, void *stream, const ngp_call_opts *opts);
NGP_API int ngp_cast_f16(const float *src, void *dst, size_t n, void *stream);
/* get_rays (nerf/utils.py:52-136) after the pixel indices are chosen: poses [B,4,4] f32 camera-to-world, intrinsics,
 * image width W; inds: pixel indices (row-major, j * W + i) [N] for all poses (inds_stride 0), [B,N] (inds_stride N) or
 * NULL (= 0..N-1, every pixel); rays_o, rays_d [B,N,3] f32, directions normalised in camera space then rotated. */
NGP_API int ngp_get_rays(const float *poses, uint32_t B, float fx, float fy, float cx, float cy, uint32_t W, const int *inds,
                 uint32_t inds_stride, uint32_t N, float *rays_o, float *rays_d, void *stream);
/* nerf/renderer.py:582-584: grid = max(grid*decay, tmp) where both >= 0; *sum_out = sum(clamp(grid, 0)) carried in DOUBLE
 * (device), so that mean = (float)(*sum_out / n) is the correctly rounded fp32 mean in any addition order */
NGP_API int ngp_density_ema(float *grid, const float *tmp, float decay, uint32_t n, double *sum_out, void *stream);
/* renderer.py:589-590: packbits with thresh = min((float)(*sum * inv_count), density_thresh) read on the device;
 * N = output bytes (N % 4 == 0) */
NGP_API int ngp_packbits_mean(const float *grid, uint32_t N, const double *sum, double inv_count, float density_thresh, uint8_t *bitfield,
                      void *stream);

/* ---------------------------------------------------------------------------------------------
 * diagnostics
 * ------------------------------------------------------------------------------------------- */
/* Runs nk tcgen05.mma (kind::f16, M=128) instructions on caller-built shared-memory byte images and
 * dumps the fp32 accumulator D[128, N] (tests/test_gpu_umma.py pins the operand-image conventions). */
NGP_API int ngp_umma_selftest(const void *a_img, uint32_t a_bytes, const void *b_img, uint32_t b_bytes, float *D, uint32_t N,
                      uint32_t nk, uint32_t idesc, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_step, uint32_t b_lbo,
                      uint32_t b_sbo, uint32_t b_step, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* NGP_B200_H */
