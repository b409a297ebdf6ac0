"""Device-resident Instant-NGP training / rendering driver ("next" row 1 of the scope table).

Restates, on top of the C-ABI ops, what /root/reference/nerf/renderer.py::run_cuda (:257-431) and
update_extra_state (:498-598) do around the hot-path kernels, with the `network_ff` topology
(/root/reference/nerf/network_ff.py:11-74: hash grid 16x2 -> FFMLP 32->64->64->16 -> trunc_exp | SH-16 + 15 geo
features + pad -> FFMLP 32->64->64->64->16 -> sigmoid), the `main_nerf -O` step (fp16 autocast + GradScaler +
Adam(lr 1e-2, betas (0.9, 0.99), eps 1e-15) + 0.1^(step/30000) decay, main_nerf.py:194-197, nerf/utils.py:960-1011)
and the upstream density-grid schedule (every 16 steps, nerf/utils.py:820-823).

B200-first differences (all below the plugin API, results equal to the autograd path through the plugin modules):
  * every buffer is allocated once at fixed capacity; a step is a fixed launch sequence with no host reads, so it is
    captured in ONE CUDA graph (the reference pays ~10 small kernels + dozens of ATen ops + 3 fresh zero-filled
    [M,.] buffers + a dense Adam + a loss.item() sync per step);
  * parameters live in one flat fp32 vector [hash table | sigma MLP | colour MLP] with an fp16 shadow of the table
    maintained by the fused Adam kernel (the reference re-casts the 49 MB table every forward under autocast);
  * gradients are one flat buffer [fp16 table | fp32 MLP] -> data-parallel training is one NCCL all-reduce per dtype.
"""
import math

import numpy as np
import torch

import ngp_abi as abi
import ffmlp as _ffmlp
import gridencoder as _gridencoder
import raymarching as _raymarching

_ST_STEP, _ST_LR, _ST_BC1, _ST_BC2, _ST_SCALE, _ST_GROWTH, _ST_FOUND_INF, _ST_LOSS, _ST_LAST_LOSS, _ST_LR0, _ST_DECAY_STEPS, _ST_DECAY, _ST_GROWTH_INTERVAL, _ST_ITER = range(14)


def _ck(rc):
    abi.check(rc)
    return rc


class NGPTrainer:
    def __init__(self, bound=1.0, num_rays=4096, device="cuda", lr=1e-2, grid_size=128, density_thresh=10.0, density_scale=1.0,
                 min_near=0.2, dt_gamma=0.0, max_steps=1024, bg_color=1.0, T_thresh=1e-4, perturb=True, loss_scale=65536.0,
                 growth_interval=2000, decay_steps=30000, world_size=1, seed=0, use_graph=True, update_interval=16,
                 num_levels=16, log2_hashmap_size=19, base_resolution=16, overlap_optimizer=False, init_samples_per_ray=512,
                 fused_nets=True):
        """loss_scale / growth_interval: torch.cuda.amp.GradScaler's defaults (init_scale 2^16, growth every 2000 clean steps, halved
        on a non-finite gradient), which is what `main_nerf -O` trains with (nerf/utils.py:328); the learning-rate decay counts
        iterations, skipped optimizer steps included, like the reference's per-iteration LambdaLR.step().
        overlap_optimizer: defer the optimizer of step n to the start of step n+1 and run it on a second stream next to
        the ray marching of step n+1 (the marcher reads rays and the occupancy bitfield, never the parameters; Adam is
        HBM-bound, the marcher instruction-bound).  Same arithmetic in the same order; the parameters a caller reads
        between two steps lag one update until flush() -- loss(), update_extra_state(), density() and render() flush or
        read around it themselves."""
        self.dev = torch.device(device)
        # fused_nets: the glue between the two MLPs (trunc_exp, SH, geo-feature shuffle, sigmoid and their backward) runs inside the
        # tcgen05 MLP kernels (ngp_nerf_sigma_forward / color_forward / color_backward); False keeps the four elementwise kernels
        # and the cin / c / grad_c / grad_cin buffers between the plain FFMLP calls -- same results bit for bit (tests).
        self.fused_nets = bool(fused_nets)
        self.merged_forward = True  # with fused_nets: both nets' forward in one launch (ngp_nerf_nets_forward); False: one launch per net
        # the warp-per-ray compositor backward writes every sample slot of every ray (zeros behind the cut-off): no zero fill per step
        self._bwd_writes_all = abi.get_option("composite_variant") != 0
        self.overlap_optimizer = bool(overlap_optimizer)
        self._pending = False
        self._feed = None
        self._side = None
        self._g_steady = None
        self.adam_blocks_per_sm = 5  # grid cap of k_adam while it shares the SMs with the marcher
        self.bound, self.N = float(bound), int(num_rays)
        self.cascade = 1 + math.ceil(math.log2(bound))
        self.H = grid_size
        self.density_thresh, self.density_scale = density_thresh, density_scale
        self.min_near, self.dt_gamma, self.max_steps = min_near, dt_gamma, max_steps
        self.bg_color, self.T_thresh, self.perturb = bg_color, T_thresh, perturb
        self.world, self.use_graph, self.update_interval = world_size, use_graph, update_interval
        self.beta1, self.beta2, self.eps = 0.9, 0.99, 1e-15
        dev = self.dev

        # ---- parameters: [hash table | sigma net | colour net], same shapes / init as the reference modules ----
        enc = _gridencoder.GridEncoder(num_levels=num_levels, log2_hashmap_size=log2_hashmap_size, base_resolution=base_resolution,
                                       desired_resolution=2048 * bound)
        self.L, self.C, self.per_level_scale, self.base_resolution = enc.num_levels, enc.level_dim, enc.per_level_scale, enc.base_resolution
        self.S = float(np.log2(enc.per_level_scale))
        self.offsets = enc.offsets.to(dev)
        self.offsets_host = np.ascontiguousarray(enc.offsets.numpy().astype(np.int32))
        self.T = int(enc.offsets[-1])
        self.enc_dim = self.L * self.C
        # the per-level-kind kernels (L % 4 == 0, level slices on 8-entry boundaries) can map [-bound, bound] -> [0, 1] themselves
        self.fuse_x01 = self.L % 4 == 0 and bool((self.offsets_host % 8 == 0).all())
        assert self.enc_dim == 32, "the network_ff topology expects a 32-wide encoding"
        sig = _ffmlp.FFMLP(32, 16, 64, 2)
        col = _ffmlp.FFMLP(32, 3, 64, 3)
        self.n16 = self.T * self.C
        self.n_sig, self.n_col = sig.num_parameters, col.num_parameters
        self.n32 = self.n_sig + self.n_col
        g = torch.Generator().manual_seed(seed)
        table0 = (torch.rand(self.T, self.C, generator=g) * 2 - 1) * 1e-4
        self.params = torch.cat([table0.reshape(-1), sig.weights.detach(), col.weights.detach()]).to(dev).contiguous()
        self.exp_avg = torch.zeros_like(self.params)
        self.exp_avg_sq = torch.zeros_like(self.params)
        self.embeddings = self.params[:self.n16].view(self.T, self.C)
        self.sigma_w = self.params[self.n16:self.n16 + self.n_sig]
        self.color_w = self.params[self.n16 + self.n_sig:]
        self.table16 = torch.empty(self.T, self.C, dtype=torch.half, device=dev)
        _ck(abi.lib.ngp_cast_f16(self.params.data_ptr(), self.table16.data_ptr(), self.n16, abi.stream()))
        self.g16 = torch.zeros(self.T, self.C, dtype=torch.half, device=dev)
        self.g32 = torch.zeros(self.n32, device=dev)
        self.g_sig, self.g_col = self.g32[:self.n_sig], self.g32[self.n_sig:]

        # ---- scene state (nerf/renderer.py:83-100) ----
        self.aabb = torch.tensor([-bound, -bound, -bound, bound, bound, bound], dtype=torch.float32, device=dev)
        self.density_grid = torch.zeros(self.cascade, self.H ** 3, device=dev)
        self.density_bitfield = torch.zeros(self.cascade * self.H ** 3 // 8, dtype=torch.uint8, device=dev)
        self.step_counter = torch.zeros(16, 2, dtype=torch.int32, device=dev)
        self.count_sum = torch.zeros(1, dtype=torch.int32, device=dev)
        self.mean_sum = torch.zeros(1, dtype=torch.float64, device=dev)  # sum of clamp(density_grid, 0), carried in double on the device
        self.mean_count, self.iter_density, self.local_step, self.global_step = -1, 0, 0, 0

        # ---- step state ----
        st = torch.zeros(16)
        st[_ST_STEP] = torch.tensor(0, dtype=torch.int32).view(torch.float32)
        st[_ST_SCALE], st[_ST_LR0], st[_ST_LR], st[_ST_DECAY_STEPS], st[_ST_DECAY] = loss_scale, lr, lr, decay_steps, 0.1
        st[_ST_GROWTH] = torch.tensor(0, dtype=torch.int32).view(torch.float32)
        st[_ST_GROWTH_INTERVAL] = torch.tensor(growth_interval, dtype=torch.int32).view(torch.float32)
        st[_ST_FOUND_INF] = torch.tensor(0, dtype=torch.int32).view(torch.float32)
        self.state = st.to(dev)

        # ---- per-ray buffers ----
        N = self.N
        f32 = dict(dtype=torch.float32, device=dev)
        self.rays_o, self.rays_d, self.gt = torch.zeros(N, 3, **f32), torch.zeros(N, 3, **f32), torch.zeros(N, 3, **f32)
        self.nears, self.fars, self.noises = torch.empty(N, **f32), torch.empty(N, **f32), torch.zeros(N, **f32)
        self.rays = torch.zeros(N, 3, dtype=torch.int32, device=dev)
        self.ws, self.depth, self.image = torch.empty(N, **f32), torch.empty(N, **f32), torch.empty(N, 3, **f32)
        self.g_image, self.g_ws, self.g_depth = torch.empty(N, 3, **f32), torch.empty(N, **f32), torch.empty(N, **f32)
        self.march_ws = torch.zeros(max(1 << 16, abi.lib.ngp_march_rays_train_workspace_bytes_ex(N, self.bound, self.max_steps)), dtype=torch.uint8, device=dev)
        self.mlp_ws = torch.zeros(1 << 17, dtype=torch.uint8, device=dev)
        self.M = 0
        self._graph = None
        # until the first density-grid interval has measured the sample count: room for 512 samples per ray (the reference
        # allocates N * max_steps and trims after a host read, raymarching.py:196-231); rays that do not fit are dropped
        self._alloc_samples(N * int(init_samples_per_ray))
        self._pool = None
        self.overlap_events = None  # set to [] to collect (side start, after all-reduce, after optimizer, after march) CUDA events per step

    # ------------------------------------------------------------------------------------------------------
    def _alloc_samples(self, M):
        M = int(max(1024, (M + 1023) // 1024 * 1024))
        if M == self.M:
            return
        dev = self.dev
        f32, f16 = dict(dtype=torch.float32, device=dev), dict(dtype=torch.half, device=dev)
        self.M = M
        self.xyzs, self.dirs, self.deltas = torch.zeros(M, 3, **f32), torch.zeros(M, 3, **f32), torch.zeros(M, 2, **f32)
        self.x01 = torch.zeros(M, 3, **f32)
        self.enc, self.h = torch.zeros(M, 32, **f16), torch.zeros(M, 16, **f16)
        Mu = 1 if self.fused_nets else M  # the matrices between the nets exist only on the unfused path
        self.cin, self.c = torch.zeros(Mu, 32, **f16), torch.zeros(Mu, 16, **f16)
        self.sigmas, self.rgbs = torch.zeros(M, **f32), torch.zeros(M, 3, **f32)
        self.g_sigmas, self.g_rgbs = torch.zeros(M, **f32), torch.zeros(M, 3, **f32)
        self.g_c, self.g_cin, self.g_h, self.g_enc = torch.zeros(Mu, 16, **f16), torch.zeros(Mu, 32, **f16), torch.zeros(M, 16, **f16), torch.zeros(M, 32, **f16)
        self._graph = None

    # ---- network pieces on raw buffers (shared by the step, the density query and the renderer) -----------
    def _opts(self, limit=None, grid=False):
        """Per-call options (ngp_call_opts): `limit` = device int32 holding the number of rows really in use (the kernels
        stop at min(M, *limit), read on the device); grid=True adds the [-bound, bound] -> [0, 1] input mapping of
        GridEncoder.forward for the grid kernels.  Passed with every call: nothing is set process-wide."""
        if limit is None and not (grid and self.fuse_x01):
            return None
        return abi.call_opts(batch_limit=limit, input_bound=self.bound if (grid and self.fuse_x01) else 0.0)

    def _encode(self, xyzs, x01, enc, M, limit=None):
        """xyzs in [-bound, bound] -> hash-grid features.  The (x + bound) / (2 bound) of GridEncoder.forward is applied by
        the grid kernels as they load the points (ngp_call_opts.input_bound) when the fast kernels serve this table;
        otherwise through the x01 buffer."""
        s = abi.stream()
        if self.fuse_x01:
            src = xyzs
        else:
            src = x01
            _ck(abi.lib.ngp_nerf_x01(xyzs.data_ptr(), self.bound, x01.data_ptr(), M, s, self._opts(limit)))
        _ck(abi.lib.ngp_grid_encode_forward(src.data_ptr(), self.table16.data_ptr(), self.offsets.data_ptr(), self.offsets_host.ctypes.data,
                                            enc.data_ptr(), M, 3, self.C, self.L, self.S, self.base_resolution, None, 0, 0, 0, abi.DTYPE_F16,
                                            abi.LAYOUT_BLC, s, self._opts(limit, grid=True)))

    def _sigma_forward(self, enc, h, M, inference, limit=None):
        fn = abi.lib.ngp_ffmlp_inference if inference else abi.lib.ngp_ffmlp_forward
        _ck(fn(enc.data_ptr(), self.sigma_w.data_ptr(), abi.DTYPE_F32, M, 32, 16, 64, 2, 0, 6, None, h.data_ptr(), abi.stream(), self._opts(limit)))

    def _color_forward(self, cin, c, M, inference, limit=None):
        fn = abi.lib.ngp_ffmlp_inference if inference else abi.lib.ngp_ffmlp_forward
        _ck(fn(cin.data_ptr(), self.color_w.data_ptr(), abi.DTYPE_F32, M, 32, 16, 64, 3, 0, 6, None, c.data_ptr(), abi.stream(), self._opts(limit)))

    def _network_forward(self, M, inference=False, limit=None):
        """xyzs, dirs -> sigmas, rgbs on the sample buffers (network_ff.py:51-74)."""
        s = abi.stream()
        self._encode(self.xyzs, self.x01, self.enc, M, limit)
        if self.fused_nets:
            self._nets_forward_fused(self.enc, self.dirs, self.h, self.sigmas, self.rgbs, M, limit)
            return
        self._sigma_forward(self.enc, self.h, M, inference, limit)
        _ck(abi.lib.ngp_nerf_mid_forward(self.h.data_ptr(), self.dirs.data_ptr(), self.density_scale, self.sigmas.data_ptr(), self.cin.data_ptr(), M, s,
                                         self._opts(limit)))
        self._color_forward(self.cin, self.c, M, inference, limit)
        _ck(abi.lib.ngp_nerf_rgb_forward(self.c.data_ptr(), self.rgbs.data_ptr(), M, s, self._opts(limit)))

    def _nets_forward_fused(self, enc, dirs, h, sigmas, rgbs, M, limit=None, keep_h=True):
        """enc, dirs -> (h,) sigmas, rgbs (network_ff.py:55-73): one launch for both nets; keep_h=False (inference) leaves h on the SM"""
        s, o = abi.stream(), self._opts(limit)
        if self.merged_forward:
            _ck(abi.lib.ngp_nerf_nets_forward(enc.data_ptr(), dirs.data_ptr(), self.sigma_w.data_ptr(), self.color_w.data_ptr(), abi.DTYPE_F32, M, 2, 3,
                                              self.density_scale, h.data_ptr() if keep_h else None, sigmas.data_ptr(), rgbs.data_ptr(), s, o))
            return
        _ck(abi.lib.ngp_nerf_sigma_forward(enc.data_ptr(), self.sigma_w.data_ptr(), abi.DTYPE_F32, M, 2, self.density_scale, h.data_ptr(),
                                           sigmas.data_ptr(), s, o))
        _ck(abi.lib.ngp_nerf_color_forward(dirs.data_ptr(), h.data_ptr(), self.color_w.data_ptr(), abi.DTYPE_F32, M, 3, rgbs.data_ptr(), s, o))

    # ---- one training step as a fixed launch sequence ---------------------------------------------------------
    def _march(self):
        """Phase A of a step: rays -> samples.  Touches neither the parameters nor the optimizer state."""
        s, N, M, lib = abi.stream(), self.N, self.M, abi.lib
        counter = self.step_counter[0]  # one slot: sample placement starts at counter[0], so it is cleared every step
        counter.zero_()
        if self.perturb:
            self.noises.uniform_()
        _ck(lib.ngp_near_far_from_aabb(self.rays_o.data_ptr(), self.rays_d.data_ptr(), self.aabb.data_ptr(), N, self.min_near,
                                       self.nears.data_ptr(), self.fars.data_ptr(), s))
        _ck(lib.ngp_march_rays_train(self.rays_o.data_ptr(), self.rays_d.data_ptr(), self.density_bitfield.data_ptr(), self.bound, self.dt_gamma,
                                     self.max_steps, N, self.cascade, self.H, M, self.nears.data_ptr(), self.fars.data_ptr(), self.xyzs.data_ptr(),
                                     self.dirs.data_ptr(), self.deltas.data_ptr(), self.rays.data_ptr(), counter.data_ptr(), self.noises.data_ptr(),
                                     self.march_ws.data_ptr(), self.march_ws.numel(), s))
        self.count_sum.add_(counter[:1])  # running sample count since the last density-grid update (-> mean_count)

    def _samples(self):
        """Phase B: network forward, compositing, loss, backward on the samples of phase A."""
        s, lib = abi.stream(), abi.lib
        counter = self.step_counter[0]
        lib.ngp_step_begin(self.state.data_ptr(), self.beta1, self.beta2, s)
        # everything per-sample below runs on min(M, counter[0]) rows, read on the device: no host sync, no padded work
        self._samples_forward_backward(counter)

    def _samples_forward_backward(self, counter):
        s, N, M, lib = abi.stream(), self.N, self.M, abi.lib
        lim = self._opts(counter)
        self._network_forward(M, limit=counter)
        _ck(lib.ngp_composite_rays_train_forward(self.sigmas.data_ptr(), self.rgbs.data_ptr(), self.deltas.data_ptr(), self.rays.data_ptr(), M, N,
                                                 self.T_thresh, self.ws.data_ptr(), self.depth.data_ptr(), self.image.data_ptr(), s))
        _ck(lib.ngp_nerf_loss(self.image.data_ptr(), self.ws.data_ptr(), self.gt.data_ptr(), self.bg_color, N, self.state[_ST_SCALE:].data_ptr(),
                              self.g_image.data_ptr(), self.g_ws.data_ptr(), self.g_depth.data_ptr(), self.state[_ST_LOSS:].data_ptr(), s))
        # ---- backward ----
        if not self._bwd_writes_all:  # the thread-per-ray compositor leaves slots behind a ray's cut-off untouched (like the reference's)
            self.g_sigmas.zero_()
            self.g_rgbs.zero_()
        _ck(lib.ngp_composite_rays_train_backward(self.g_ws.data_ptr(), self.g_depth.data_ptr(), self.g_image.data_ptr(), self.sigmas.data_ptr(),
                                                  self.rgbs.data_ptr(), self.deltas.data_ptr(), self.rays.data_ptr(), self.ws.data_ptr(),
                                                  self.depth.data_ptr(), self.image.data_ptr(), M, N, self.T_thresh, self.g_sigmas.data_ptr(),
                                                  self.g_rgbs.data_ptr(), s))
        if self.fused_nets:
            _ck(lib.ngp_nerf_color_backward(self.g_rgbs.data_ptr(), self.rgbs.data_ptr(), self.g_sigmas.data_ptr(), self.dirs.data_ptr(), self.h.data_ptr(),
                                            self.color_w.data_ptr(), abi.DTYPE_F32, M, 3, self.density_scale, self.g_h.data_ptr(), self.g_col.data_ptr(),
                                            abi.DTYPE_F32, self.mlp_ws.data_ptr(), self.mlp_ws.numel(), s, lim))
        else:
            _ck(lib.ngp_nerf_rgb_backward(self.g_rgbs.data_ptr(), self.rgbs.data_ptr(), self.g_c.data_ptr(), M, s, lim))
            _ck(lib.ngp_ffmlp_backward(self.g_c.data_ptr(), self.cin.data_ptr(), self.color_w.data_ptr(), abi.DTYPE_F32, None, M, 32, 16, 64, 3, 0, 6, 1,
                                       None, self.g_cin.data_ptr(), self.g_col.data_ptr(), abi.DTYPE_F32, self.mlp_ws.data_ptr(), self.mlp_ws.numel(), s, lim))
            _ck(lib.ngp_nerf_mid_backward(self.g_sigmas.data_ptr(), self.h.data_ptr(), self.g_cin.data_ptr(), self.density_scale, self.g_h.data_ptr(), M, s, lim))
        _ck(lib.ngp_ffmlp_backward(self.g_h.data_ptr(), self.enc.data_ptr(), self.sigma_w.data_ptr(), abi.DTYPE_F32, None, M, 32, 16, 64, 2, 0, 6, 1,
                                   None, self.g_enc.data_ptr(), self.g_sig.data_ptr(), abi.DTYPE_F32, self.mlp_ws.data_ptr(), self.mlp_ws.numel(), s, lim))
        # table gradient: g16 was cleared by the previous optimizer step (k_adam consumes and zeroes it)
        _ck(lib.ngp_grid_encode_backward(self.g_enc.data_ptr(), (self.xyzs if self.fuse_x01 else self.x01).data_ptr(), self.table16.data_ptr(),
                                         self.offsets.data_ptr(), self.offsets_host.ctypes.data, self.g16.data_ptr(), M, 3, self.C, self.L, self.S,
                                         self.base_resolution, None, None, 0, 0, 0, abi.DTYPE_F16, abi.LAYOUT_BLC, s, self._opts(counter, grid=True)))

    def _optimizer_step(self):
        s, lib = abi.stream(), abi.lib
        _ck(lib.ngp_found_inf(self.g16.data_ptr(), self.n16, self.g32.data_ptr(), self.n32, self.state.data_ptr(), s))
        cap = abi.call_opts(grid_cap_blocks_per_sm=self.adam_blocks_per_sm) if self.overlap_optimizer else None
        _ck(lib.ngp_adam_step(self.params.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.g16.data_ptr(), self.g32.data_ptr(),
                              self.table16.data_ptr(), self.n16, self.n32, self.state.data_ptr(), self.beta1, self.beta2, self.eps, float(self.world), 0.0, s, cap))
        _ck(lib.ngp_step_end(self.state.data_ptr(), s))

    def _allreduce(self):
        from .parallel import allreduce_gradients
        allreduce_gradients(self.g16, self.g32)

    def _forward_backward(self):
        self._march()
        self._samples()

    def _step_eager(self):
        self._forward_backward()
        if self.world > 1:
            self._allreduce()
        self._optimizer_step()

    # ---- deferred optimizer (overlap_optimizer=True) -----------------------------------------------------------------
    def _run(self, what):
        """one phase, as a graph replay when captured"""
        if self.use_graph and self._graph:
            {"march": self._g_a, "samples": self._g_b, "opt": self._g_opt}[what].replay()
        else:
            {"march": self._march, "samples": self._samples, "opt": self._optimizer_step}[what]()

    def flush(self):
        """Apply the deferred optimizer step, if any, on the current stream."""
        if self._pending:
            if self.world > 1:
                self._allreduce()
            self._run("opt")
            self._pending = False

    def _step_overlapped(self):
        main = torch.cuda.current_stream()
        if self._pending and self.use_graph and self._graph and self._g_steady is not None:
            self._g_steady.replay()
            return
        if self._pending:
            if self._side is None:
                self._side = torch.cuda.Stream()
            self._side.wait_stream(main)  # the gradients of the previous step
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)] if self.overlap_events is not None else None
            self._run("march")          # submitted first: its blocks take their share of every SM before Adam's do
            if ev:
                ev[3].record(main)
            with torch.cuda.stream(self._side):
                if ev:
                    ev[0].record(self._side)
                if self.world > 1:
                    self._allreduce()
                if ev:
                    ev[1].record(self._side)
                self._run("opt")
                if ev:
                    ev[2].record(self._side)
                    self.overlap_events.append(ev)
            self._pending = False
            main.wait_stream(self._side)
        else:
            self._run("march")
        self._run("samples")
        self._pending = True

    def _capture(self):
        """Warm up eagerly on a side stream (lazy kernel attributes, NCCL communicators), then capture the launch sequence.
        With several ranks the all-reduce stays outside the graphs: [forward+backward graph] -> NCCL -> [optimizer graph]."""
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            self._step_eager()
        torch.cuda.current_stream().wait_stream(stream)
        torch.cuda.synchronize()
        self._g_fb = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g_fb):
            self._forward_backward()
        if self.overlap_optimizer:  # the same launches as two graphs, so the optimizer can run beside the first
            self._g_a = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._g_a):
                self._march()
            self._g_b = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._g_b):
                self._samples()
            self._g_steady = None
            if self.world == 1:  # one graph for the steady state: {optimizer(n-1) | march(n)} -> samples(n); one launch per step
                if self._side is None:
                    self._side = torch.cuda.Stream()
                self._g_steady = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._g_steady):
                    cur = torch.cuda.current_stream()
                    self._side.wait_stream(cur)
                    self._march()
                    with torch.cuda.stream(self._side):
                        self._optimizer_step()
                    cur.wait_stream(self._side)
                    self._samples()
        self._g_opt = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g_opt):
            self._optimizer_step()
        self._graph = True
        return 1  # the warm-up was a real optimisation step

    def train_step(self, rays_o=None, rays_d=None, gt=None):
        """One optimisation step.  rays_o / rays_d / gt: [N,3] tensors (device, or pinned host -> async H2D); None keeps
        whatever is in the ray buffers (e.g. filled by next_pool_batch)."""
        if rays_o is not None:  # host batches travel on a side stream while the previous step computes (nerf/feed.py)
            if self._feed is None:
                from .feed import HostFeed
                self._feed = HostFeed(self.dev)
            self._feed.stage((self.rays_o, self.rays_d, self.gt), (rays_o, rays_d, gt))
        if self.update_interval > 0 and self.global_step % self.update_interval == 0:
            self.update_extra_state()
        if self.use_graph and self._graph is None:
            self.flush()
            self._capture()
        elif self.overlap_optimizer:
            self._step_overlapped()
        elif self.use_graph:
            self._g_fb.replay()
            if self.world > 1:
                self._allreduce()
            self._g_opt.replay()
        else:
            self._step_eager()
        self.local_step += 1
        self.global_step += 1

    def loss(self):
        """Loss of the last finished step (device -> host read).  With a deferred optimizer step pending, the value is
        still in the accumulator slot (k_step_end has not copied it yet): read it there instead of forcing the update."""
        return float(self.state[_ST_LOSS if self._pending else _ST_LAST_LOSS].item())

    def loss_async(self):
        """The same read without stalling the launch queue: enqueues the device -> host copy of the last step's loss into pinned
        memory on the current stream (it lands after that step's kernels and before the next step clears the accumulator) and
        returns a function that waits for THAT copy only and gives the value.  A training loop calls it after every step and
        collects the value one step later, so the host keeps launching while the GPU works."""
        ring = self.__dict__.setdefault("_loss_ring", {"i": 0, "buf": torch.zeros(8, dtype=torch.float32).pin_memory(),
                                                        "ev": [torch.cuda.Event() for _ in range(8)]})
        i = ring["i"] = (ring["i"] + 1) % 8
        slot = _ST_LOSS if self._pending else _ST_LAST_LOSS
        ring["buf"][i:i + 1].copy_(self.state[slot:slot + 1], non_blocking=True)
        ev = ring["ev"][i]
        ev.record()

        def get():
            ev.synchronize()
            return float(ring["buf"][i])
        return get

    # ---- synthetic ray pool (device resident inputs for the kernel-side benchmark) ----------------------------
    def set_pool(self, rays_o, rays_d, gt):
        self._pool = (rays_o.contiguous(), rays_d.contiguous(), gt.contiguous())
        self._cursor = 0

    def next_pool_batch(self):
        o, d, g = self._pool
        P, N = o.shape[0], self.N
        if self._cursor + N > P:
            self._cursor = 0
        sl = slice(self._cursor, self._cursor + N)
        self.rays_o.copy_(o[sl]); self.rays_d.copy_(d[sl]); self.gt.copy_(g[sl])
        self._cursor += N

    # ---- density grid maintenance (nerf/renderer.py:498-598), eager, every `update_interval` steps ---------------
    @torch.no_grad()
    def density(self, xyz):
        """sigma(x) * density_scale through the encoder + sigma net in inference mode; xyz [n,3] in [-bound, bound]."""
        self.flush()
        n = xyz.shape[0]
        out = torch.empty(n, device=self.dev)
        chunk = 1 << 21
        x01 = torch.empty(min(n, chunk), 3, device=self.dev)
        enc = torch.empty(min(n, chunk), 32, dtype=torch.half, device=self.dev)
        h = torch.empty(min(n, chunk), 16, dtype=torch.half, device=self.dev)
        for a in range(0, n, chunk):
            b = min(n, a + chunk)
            xs = xyz[a:b].contiguous()
            self._encode(xs, x01, enc, b - a)
            self._sigma_forward(enc, h, b - a, True)
            out[a:b] = torch.exp(h[:b - a, 0].float()) * self.density_scale
        return out

    @torch.no_grad()
    def update_extra_state(self, decay=0.95):
        """nerf/renderer.py:498-598.  Cell sampling follows the reference call for call (same torch RNG calls in the same
        order, so a seeded run probes the same points): full sweep of all H^3 cells for the first 16 updates (:510-538, one
        S = H block), then per cascade H^3/4 uniform cells + H^3/4 cells re-drawn from the occupied ones (:546-574).  The
        EMA (:583), mean of clamp(min=0) (:584), min(mean, thresh) (:589) and packbits (:590) are two kernels with the mean
        kept on the device; step-counter bookkeeping (:593-596) uses a device-side running sum."""
        self.flush()
        H, dev = self.H, self.dev
        tmp = -torch.ones_like(self.density_grid)

        def probe(cas, coords, indices):
            xyzs = 2 * coords.float() / (H - 1) - 1
            bound = min(2 ** cas, self.bound)
            hgs = bound / H
            cas_xyzs = xyzs * (bound - hgs)
            cas_xyzs += (torch.rand_like(cas_xyzs) * 2 - 1) * hgs
            tmp[cas, indices] = self.density(cas_xyzs)

        if self.iter_density < 16:  # full sweep
            if not hasattr(self, "_all_coords"):
                r = torch.arange(H, dtype=torch.int32, device=dev)
                xx, yy, zz = torch.meshgrid(r, r, r, indexing="ij")
                self._all_coords = torch.stack([xx.reshape(-1), yy.reshape(-1), zz.reshape(-1)], -1).contiguous()
                self._all_indices = _raymarching.morton3D(self._all_coords).long()
            for cas in range(self.cascade):
                probe(cas, self._all_coords, self._all_indices)
        else:  # partial: H^3/4 uniform cells + H^3/4 cells resampled from the occupied ones, per cascade
            n = H ** 3 // 4
            for cas in range(self.cascade):
                coords = torch.randint(0, H, (n, 3), device=dev)  # int64 like the reference's call (:549); morton3D takes .int()
                indices = _raymarching.morton3D(coords.int()).long()
                occ = torch.nonzero(self.density_grid[cas] > 0).squeeze(-1)
                if occ.shape[0] > 0:  # (the reference's randint(0, 0) raises on an empty grid)
                    occ = occ[torch.randint(0, occ.shape[0], [n], dtype=torch.long, device=dev)]
                    occ_coords = _raymarching.morton3D_invert(occ.int())
                    indices = torch.cat([indices, occ])
                    coords = torch.cat([coords, occ_coords])
                probe(cas, coords, indices)
        self._density_ema_packbits(tmp, decay)
        self.iter_density += 1
        # sample budget for the next interval (the one host read of the schedule, nerf/renderer.py:593-596)
        if self.local_step > 0:
            self.mean_count = int(self.count_sum.item() / self.local_step)
            need = int(self.mean_count * 1.15) + 4096  # capacity only: the kernels stop at the device-side count
            if need > self.M or need < 0.7 * self.M:
                self._alloc_samples(need)
        self.count_sum.zero_()
        self.local_step = 0

    def _density_ema_packbits(self, tmp, decay):
        """renderer.py:582-590 on the device: grid = max(grid * decay, tmp) where both >= 0; mean of clamp(grid, 0);
        bitfield = grid > min(mean, density_thresh)."""
        n_cells = self.density_grid.numel()
        _ck(abi.lib.ngp_density_ema(self.density_grid.data_ptr(), tmp.data_ptr(), decay, n_cells, self.mean_sum.data_ptr(), abi.stream()))
        _ck(abi.lib.ngp_packbits_mean(self.density_grid.data_ptr(), n_cells // 8, self.mean_sum.data_ptr(), 1.0 / n_cells, self.density_thresh,
                                      self.density_bitfield.data_ptr(), abi.stream()))

    @torch.no_grad()
    def mark_untrained_grid(self, poses, intrinsic):
        """nerf/renderer.py:433-496: cells outside every training camera's frustum are set to -1 and stay unoccupied.
        Returns the number of cells marked (the figure the reference prints)."""
        self.flush()
        count = _raymarching.mark_untrained_grid(poses, intrinsic, self.density_grid, self.bound, return_count=True)
        return int((count == 0).sum().item())

    def mean_density(self):
        """mean(clamp(density_grid, 0)) of the last update as the fp32 value the bitfield threshold used (renderer.py:584)."""
        return float(np.float32(float(self.mean_sum.item()) / self.density_grid.numel()))

    # ---- checkpoints in the reference's layout (nerf/utils.py:1175-1311, state-dict names of network_ff.NeRFNetwork) ----
    def state_dict(self):
        """Model state with the names the reference's NeRFNetwork(--ff) registers: encoder.embeddings / encoder.offsets
        (gridencoder/grid.py:787-798), sigma_net.weights / color_net.weights (ffmlp.py:121-122), and the renderer buffers
        aabb_train, aabb_infer, density_grid, density_bitfield, step_counter (renderer.py:85-100)."""
        self.flush()
        return {
            "encoder.embeddings": self.embeddings.detach().clone(), "encoder.offsets": self.offsets.detach().clone(),
            "sigma_net.weights": self.sigma_w.detach().clone(), "color_net.weights": self.color_w.detach().clone(),
            "aabb_train": self.aabb.clone(), "aabb_infer": self.aabb.clone(), "density_grid": self.density_grid.clone(),
            "density_bitfield": self.density_bitfield.clone(), "step_counter": self.step_counter.clone(),
        }

    @torch.no_grad()
    def load_state_dict(self, sd, strict=True):
        """Inverse of state_dict(); also accepts a state dict written by the reference.  Returns (missing, unexpected)."""
        self.flush()
        mine = {"encoder.embeddings": self.embeddings, "sigma_net.weights": self.sigma_w, "color_net.weights": self.color_w,
                "density_grid": self.density_grid, "density_bitfield": self.density_bitfield, "step_counter": self.step_counter}
        known = set(mine) | {"encoder.offsets", "aabb_train", "aabb_infer"}
        missing = [k for k in mine if k not in sd]
        unexpected = [k for k in sd if k not in known]
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing {missing}, unexpected {unexpected}")
        if "encoder.offsets" in sd and not torch.equal(sd["encoder.offsets"].cpu().int(), self.offsets.cpu().int()):
            raise RuntimeError("load_state_dict: the checkpoint's hash-grid level offsets differ from this trainer's configuration")
        for k, dst in mine.items():
            if k in sd:
                src = sd[k]
                if src.numel() != dst.numel():
                    raise RuntimeError(f"load_state_dict: size mismatch for {k}: {tuple(src.shape)} vs {tuple(dst.shape)}")
                dst.copy_(src.to(dst.device, dst.dtype).view_as(dst))
        _ck(abi.lib.ngp_cast_f16(self.params.data_ptr(), self.table16.data_ptr(), self.n16, abi.stream()))  # fp16 shadow of the table
        return missing, unexpected

    def save_checkpoint(self, path, full=True):
        """{'epoch', 'global_step', 'stats', 'mean_count', 'mean_density', 'model'} as Trainer.save_checkpoint writes them
        (utils.py:1180-1199; its load_checkpoint indexes 'stats', 'epoch' and 'global_step' unconditionally, :1270-1272, so
        they are always present: the epoch counter and the stats lists are the reference Trainer's bookkeeping, kept here
        only so that its loader accepts the file); full adds this trainer's optimizer state (Adam moments, step / loss-scale
        state) under 'optimizer_b200'."""
        state = {"epoch": getattr(self, "epoch", 0), "global_step": self.global_step,
                 "stats": getattr(self, "stats", {"loss": [], "valid_loss": [], "results": [], "checkpoints": [], "best_result": None}),
                 "mean_count": self.mean_count, "mean_density": self.mean_density(), "model": self.state_dict()}
        if full:
            state["optimizer_b200"] = {"exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(), "state": self.state.clone(),
                                       "iter_density": self.iter_density, "local_step": self.local_step, "count_sum": self.count_sum.clone()}
        torch.save(state, path)

    def load_checkpoint(self, path, model_only=False):
        ck = torch.load(path, map_location=self.dev)
        if "model" not in ck:  # a bare state dict (utils.py:1249-1252)
            self.load_state_dict(ck, strict=False)
            return
        self.load_state_dict(ck["model"], strict=False)
        if "mean_count" in ck:
            self.mean_count = ck["mean_count"]
        if "mean_density" in ck:  # utils.py:1259-1260; kept as the running sum the device-side threshold reads
            self.mean_sum.fill_(float(ck["mean_density"]) * self.density_grid.numel())
        else:
            self.mean_sum.copy_(self.density_grid.double().clamp(min=0).sum().view(1))
        self.count_sum.zero_()  # sample counts of the interval that was running when the file was written are not in it
        self.local_step = 0
        self.global_step = ck.get("global_step", self.global_step)
        self.epoch, self.stats = ck.get("epoch", 0), ck.get("stats", getattr(self, "stats", None) or {})
        if not model_only and "optimizer_b200" in ck:
            o = ck["optimizer_b200"]
            self.exp_avg.copy_(o["exp_avg"]); self.exp_avg_sq.copy_(o["exp_avg_sq"]); self.state.copy_(o["state"])
            self.iter_density = o["iter_density"]
            if "count_sum" in o:  # the running sample count and the number of steps it covers travel together
                self.count_sum.copy_(o["count_sum"])
                self.local_step = o["local_step"]

    # ---- inference rendering (nerf/renderer.py:346-431, the `not self.training` branch of run_cuda) ----------------
    def _render_state(self, N, rows):
        """Buffers of the render loop for N rays and at most `rows` samples per iteration, allocated once and kept (a frame
        sequence renders the same N again and again), plus the CUDA graphs of its iteration groups."""
        cache = self.__dict__.setdefault("_rcache", {})
        rs = cache.get((N, rows))
        if rs is not None:
            return rs
        if len(cache) >= 4:
            cache.pop(next(iter(cache)))
        dev, lib = self.dev, abi.lib
        f32, f16 = dict(dtype=torch.float32, device=dev), dict(dtype=torch.half, device=dev)
        cap = rows + 128  # n_alive * n_step <= rows by construction
        rs = dict(
            rays_o=torch.empty(N, 3, **f32), rays_d=torch.empty(N, 3, **f32), nears=torch.empty(N, **f32), fars=torch.empty(N, **f32),
            ws=torch.empty(N, **f32), depth=torch.empty(N, **f32), image=torch.empty(N, 3, **f32), rays_t=torch.empty(N, **f32),
            alive=[torch.empty(N, dtype=torch.int32, device=dev), torch.empty(N, dtype=torch.int32, device=dev)],
            iota=torch.arange(N, dtype=torch.int32, device=dev), cnt2=torch.zeros(2, dtype=torch.int32, device=dev),
            work=torch.zeros(2, dtype=torch.int64, device=dev),
            xyzs=torch.zeros(cap, 3, **f32), dirs=torch.zeros(cap, 3, **f32), deltas=torch.zeros(cap, 2, **f32), x01=torch.empty(cap, 3, **f32),
            enc=torch.empty(cap, 32, **f16), h=torch.empty(cap, 16, **f16), cin=torch.empty(1 if self.fused_nets else cap, 32, **f16),
            c=torch.empty(1 if self.fused_nets else cap, 16, **f16),
            sigmas=torch.empty(cap, **f32), rgbs=torch.empty(cap, 3, **f32),
            cws=torch.empty(max(1 << 16, lib.ngp_compact_alive_workspace_bytes(N)), dtype=torch.uint8, device=dev),
            noises=torch.zeros(N, **f32), zeros_n=torch.zeros(N, **f32), graphs={}, table=None)
        rs["rows_dev"], rs["n_dev"] = rs["cnt2"][0:1], rs["cnt2"][1:2]  # two views of one buffer: one add keeps the frame totals
        cache[(N, rows)] = rs
        return rs

    def _render_iteration(self, rs, n_alive, n_step, cur, noises, T_thresh):
        """One loop iteration (renderer.py:380-417) for at most n_alive rays (the kernels clamp to the device count)."""
        lib, s = abi.lib, abi.stream()
        M = n_alive * n_step
        n_dev = rs["n_dev"]
        rs["deltas"][:M].zero_()  # a zero delta ends a ray's samples in composite_rays (the reference allocates zeroed buffers every iteration)
        torch.mul(n_dev, n_step, out=rs["rows_dev"])  # sample rows really in use: bounds the shading kernels on the device
        rs["work"].add_(rs["cnt2"])  # running totals of the frame: sample rows shaded, ray-iterations marched (for the roofline)
        _ck(lib.ngp_march_rays(n_alive, n_dev.data_ptr(), n_step, rs["alive"][cur].data_ptr(), rs["rays_t"].data_ptr(), rs["rays_o"].data_ptr(),
                               rs["rays_d"].data_ptr(), self.bound, self.dt_gamma, self.max_steps, self.cascade, self.H,
                               self.density_bitfield.data_ptr(), rs["nears"].data_ptr(), rs["fars"].data_ptr(), rs["xyzs"].data_ptr(),
                               rs["dirs"].data_ptr(), rs["deltas"].data_ptr(), noises.data_ptr(), s))
        # shade the samples (inference kernels: no activations kept); rows_dev bounds every kernel on the device
        lim, rows_dev = self._opts(rs["rows_dev"]), rs["rows_dev"]
        self._encode(rs["xyzs"], rs["x01"], rs["enc"], M, rows_dev)
        if self.fused_nets:
            self._nets_forward_fused(rs["enc"], rs["dirs"], rs["h"], rs["sigmas"], rs["rgbs"], M, rows_dev, keep_h=False)
        else:
            _ck(lib.ngp_ffmlp_inference(rs["enc"].data_ptr(), self.sigma_w.data_ptr(), abi.DTYPE_F32, M, 32, 16, 64, 2, 0, 6, None, rs["h"].data_ptr(), s, lim))
            _ck(lib.ngp_nerf_mid_forward(rs["h"].data_ptr(), rs["dirs"].data_ptr(), self.density_scale, rs["sigmas"].data_ptr(), rs["cin"].data_ptr(), M, s, lim))
            _ck(lib.ngp_ffmlp_inference(rs["cin"].data_ptr(), self.color_w.data_ptr(), abi.DTYPE_F32, M, 32, 16, 64, 3, 0, 6, None, rs["c"].data_ptr(), s, lim))
            _ck(lib.ngp_nerf_rgb_forward(rs["c"].data_ptr(), rs["rgbs"].data_ptr(), M, s, lim))
        _ck(lib.ngp_composite_rays(n_alive, n_dev.data_ptr(), n_step, T_thresh, rs["alive"][cur].data_ptr(), rs["rays_t"].data_ptr(),
                                   rs["sigmas"].data_ptr(), rs["rgbs"].data_ptr(), rs["deltas"].data_ptr(), rs["ws"].data_ptr(), rs["depth"].data_ptr(),
                                   rs["image"].data_ptr(), s))
        _ck(lib.ngp_compact_alive(rs["alive"][cur].data_ptr(), n_alive, n_dev.data_ptr(), rs["alive"][cur ^ 1].data_ptr(), n_dev.data_ptr(),
                                  rs["cws"].data_ptr(), rs["cws"].numel(), s))

    @torch.no_grad()
    def render(self, rays_o, rays_d, perturb=False, T_thresh=1e-4, max_iters=None, sync_every=4, use_graph=True, n_step_range=(2, 64)):
        """Image [N,3], depth [N] for the given rays (the rows of a frame one rank owns when sharded, nerf/parallel.py).
        The loop of the reference: march up to n_step samples for every ray still alive, shade them, composite in place,
        drop the finished rays -- on buffers kept per ray count, with the alive list compacted on the device
        (raymarching.compact_alive).  The alive count is read back only every `sync_every` iterations: in between, the
        launches are sized by an upper bound of the count (it never grows) and the kernels clamp to the count on the
        device (n_alive_dev / ngp_call_opts.batch_limit), so the host does not wait on the GPU 200 times per frame.  With
        use_graph the bound is rounded up a 2^(1/8) ladder and each group of `sync_every` iterations with the same
        (bound, n_step) is one CUDA graph, captured on first use and replayed for every later frame of the same size.
        A different cut of a ray's samples into iterations moves the points where the march resumes from the compositor's
        running t, i.e. sample positions in their last bits -- as the reference's own n_alive-driven n_step does; images of
        different schedules agree to the tolerance of the fp16 network (tests: 2e-3), each op stays bit-exact.
        n_step_range: the reference marches n_step = clamp(N // n_alive, 1, 8) samples per ray and iteration, which makes
        >= 128 iterations for a ray that crosses the whole box and ~200 per frame, each a dozen launches whose fixed cost
        dominates once few rays are left.  (lo, hi) = (2, 64) with room for lo * N samples cuts that to a few dozen
        iterations; the only cost is the < n_step samples a ray computes past its own end.
        sync_every=1, use_graph=False, n_step_range=(1, 8) is the reference's schedule."""
        self.flush()
        lib, s, dev = abi.lib, abi.stream(), self.dev
        rays_o, rays_d = rays_o.contiguous().view(-1, 3), rays_d.contiguous().view(-1, 3)
        N = rays_o.shape[0]
        lo, hi = max(1, int(n_step_range[0])), max(1, int(n_step_range[1]))
        rows = lo * N
        rs = self._render_state(N, rows)
        if rs["table"] is not self.table16 or rs.get("M_train") != self.M:  # graphs hold the parameter pointers of their capture
            rs["graphs"].clear()
            rs["table"], rs["M_train"] = self.table16, self.M
        rs["rays_o"].copy_(rays_o, non_blocking=True)
        rs["rays_d"].copy_(rays_d, non_blocking=True)
        nears, fars = rs["nears"], rs["fars"]
        _ck(lib.ngp_near_far_from_aabb(rs["rays_o"].data_ptr(), rs["rays_d"].data_ptr(), self.aabb.data_ptr(), N, self.min_near, nears.data_ptr(),
                                       fars.data_ptr(), s))
        rs["ws"].zero_(); rs["depth"].zero_(); rs["image"].zero_(); rs["work"].zero_()
        rs["alive"][0].copy_(rs["iota"])
        rs["n_dev"].fill_(N)
        rs["rays_t"].copy_(nears)
        if perturb:
            rs["noises"].uniform_()
        else:
            rs["noises"].zero_()
        sync_every = max(1, int(sync_every))
        group = sync_every if sync_every % 2 == 0 else 1  # a graph must end on the alive buffer it started on
        n_alive, step, cur, it = N, 0, 0, 0  # n_alive: the host's view of the count (possibly stale, never too small)
        while step < self.max_steps and n_alive > 0:
            n_step = max(min(rows // n_alive, hi), lo)
            left = (self.max_steps - step + n_step - 1) // n_step  # iterations the reference's loop bound still allows
            if max_iters is not None:
                left = min(left, max_iters - it)
            if use_graph and it > 0 and group > 1 and left >= group:
                key = (n_alive, n_step, cur, float(T_thresh))
                g = rs["graphs"].get(key)
                if g is None:
                    side = torch.cuda.Stream()
                    side.wait_stream(torch.cuda.current_stream())
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=side):
                        for k in range(group):
                            self._render_iteration(rs, n_alive, n_step, cur ^ (k & 1), rs["zeros_n"], T_thresh)
                    torch.cuda.current_stream().wait_stream(side)
                    rs["graphs"][key] = g
                g.replay()
                done = group
            else:
                self._render_iteration(rs, n_alive, n_step, cur, rs["noises"] if step == 0 else rs["zeros_n"], T_thresh)
                cur ^= 1
                done = 1
            step += n_step * done
            it += done
            if it % sync_every == 0 or done > 1:
                actual = int(rs["n_dev"].item())
                n_alive = actual
                if use_graph and actual > 0:  # round up a 2^(1/8) ladder so that (bound, n_step) repeats from frame to frame
                    j = int(math.floor(8 * math.log2(N / actual) + 1e-9)) if actual < N else 0  # largest j with N * 2^(-j/8) >= actual
                    n_alive = max(actual, min(N, int(math.ceil(N * 2 ** (-j / 8)))))
            if max_iters is not None and it >= max_iters:
                break
        image = rs["image"] + (1 - rs["ws"]).unsqueeze(-1) * self.bg_color
        depth = torch.clamp(rs["depth"] - nears, min=0) / (fars - nears)
        self.last_render_iters = it
        self.last_render_work = rs["work"]  # device tensor [sample rows shaded, ray-iterations marched] of this frame
        return image, depth
