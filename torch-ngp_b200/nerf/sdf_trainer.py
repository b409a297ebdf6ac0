"""Device-resident hash-grid SDF training step (BASELINE.json configs[4]; /root/reference/main_sdf.py --fp16 --ff).

Network: sdf/netowrk_ff.py:9-50 -- hash grid (16 x 2, T = 2^19, desired resolution 2048) -> FFMLP 32 -> 64 x 3 -> 1.
Step: sdf/utils.py:217-225 (pred = model(X); mape_loss, loss.py:7-15) under fp16 autocast + GradScaler,
Adam(lr 1e-4, betas (0.9, 0.99), eps 1e-15, weight_decay 1e-6 on the MLP; main_sdf.py:51-54).
Same construction as nerf/trainer.py: flat parameters, fp16 table shadow, fixed launch sequence in two CUDA graphs.
The EMA of the weights (torch_ema, evaluation only) and the per-epoch StepLR are host-side and out of scope."""
import numpy as np
import torch

import ngp_abi as abi
import ffmlp as _ffmlp
import gridencoder as _gridencoder

from .trainer import _ST_LAST_LOSS, _ST_LOSS, _ST_SCALE, _ST_LR, _ST_LR0, _ST_DECAY_STEPS, _ST_DECAY, _ST_GROWTH_INTERVAL, _ck


class SDFTrainer:
    def __init__(self, batch=1 << 20, device="cuda", lr=1e-4, weight_decay=1e-6, loss_scale=65536.0, growth_interval=2000, world_size=1, seed=0,
                 use_graph=True, num_layers=3, hidden_dim=64, log2_hashmap_size=19, grad_chunks=None, bin_bits=7, coarse_smem=None):
        self.dev = dev = torch.device(device)
        self.B, self.world, self.use_graph = int(batch), world_size, use_graph
        self.beta1, self.beta2, self.eps, self.wd = 0.9, 0.99, 1e-15, weight_decay
        self.nl, self.hid = num_layers, hidden_dim
        # Table-gradient pipeline (world_size > 1): the grid backward runs as `grad_chunks` level ranges, finest first, and a
        # side stream all-reduces the range that is done while the next one is still being accumulated -- the step has no
        # marcher to hide the collective behind (nerf/trainer.py), so only the last, smallest range (the dense coarse
        # levels: 13 % of the table) stays exposed.  grad_chunks = 1: one backward call, then the two all-reduces.
        self.grad_chunks = int(grad_chunks) if grad_chunks is not None else (4 if world_size > 1 else 1)
        enc = _gridencoder.GridEncoder(num_levels=16, log2_hashmap_size=log2_hashmap_size, desired_resolution=2048)
        self.L, self.C, self.base_resolution = enc.num_levels, enc.level_dim, enc.base_resolution
        self.S = float(np.log2(enc.per_level_scale))
        self.per_level_scale = enc.per_level_scale
        self.offsets = enc.offsets.to(dev)
        self.offsets_host = np.ascontiguousarray(enc.offsets.numpy().astype(np.int32))
        self.T = int(enc.offsets[-1])
        net = _ffmlp.FFMLP(32, 1, hidden_dim, num_layers)
        self.n16, self.n32 = self.T * self.C, net.num_parameters
        g = torch.Generator().manual_seed(seed)
        table0 = (torch.rand(self.T, self.C, generator=g) * 2 - 1) * 1e-4
        self.params = torch.cat([table0.reshape(-1), net.weights.detach()]).to(dev).contiguous()
        self.exp_avg, self.exp_avg_sq = torch.zeros_like(self.params), torch.zeros_like(self.params)
        self.embeddings = self.params[:self.n16].view(self.T, self.C)
        self.mlp_w = self.params[self.n16:]
        self.table16 = torch.empty(self.T, self.C, dtype=torch.half, device=dev)
        _ck(abi.lib.ngp_cast_f16(self.params.data_ptr(), self.table16.data_ptr(), self.n16, abi.stream()))
        self.g16 = torch.zeros(self.T, self.C, dtype=torch.half, device=dev)
        self.g32 = torch.zeros(self.n32, device=dev)
        st = torch.zeros(16)
        i32 = lambda v: torch.tensor(v, dtype=torch.int32).view(torch.float32)
        st[0], st[5], st[6], st[_ST_GROWTH_INTERVAL] = i32(0), i32(0), i32(0), i32(growth_interval)
        st[_ST_SCALE], st[_ST_LR0], st[_ST_LR], st[_ST_DECAY_STEPS], st[_ST_DECAY] = loss_scale, lr, lr, 1e30, 1.0  # constant lr inside an epoch
        self.state = st.to(dev)
        B = self.B
        f16 = dict(dtype=torch.half, device=dev)
        self.points, self.sdfs = torch.zeros(B, 3, device=dev), torch.zeros(B, device=dev)
        self.x01 = torch.empty(B, 3, device=dev)
        self.enc, self.h = torch.empty(B, 32, **f16), torch.empty(B, 16, **f16)
        self.g_h, self.g_enc = torch.empty(B, 16, **f16), torch.empty(B, 32, **f16)
        self.mlp_ws = torch.zeros(1 << 17, dtype=torch.uint8, device=dev)
        # bin_bits in {5, 6, 7}: the batch is re-ordered by the Morton code of a (2^bits)^3 grid at the start of the step
        # (ngp_bin_points; the loss is a mean, the order of its points is free) so that the points of a cell sit in
        # neighbouring lanes: the table-gradient kernel merges them in the warp instead of in the L2 atomic unit.  0 = keep
        # the order of the batch (then the coarse levels go through shared memory, NGP_OPT_GRID_BWD_SMEM_COARSE).
        self.bin_bits = int(bin_bits)
        self.coarse_smem = (self.bin_bits == 0) if coarse_smem is None else bool(coarse_smem)
        if self.bin_bits:
            self.pts, self.tgt = torch.empty_like(self.points), torch.empty_like(self.sdfs)
            self.bin_ws = torch.zeros(int(abi.lib.ngp_bin_points_workspace(B, self.bin_bits)), dtype=torch.uint8, device=dev)
        else:
            self.pts, self.tgt = self.points, self.sdfs
        self._graph = None
        self.global_step = 0

    def _level_ranges(self):
        """[(level_begin, level_end)] of the chunked backward, finest levels first (multiples of 4)."""
        n = max(1, min(self.grad_chunks, self.L // 4))
        per = -(-(self.L // 4) // n) * 4
        cuts = list(range(0, self.L, per)) + [self.L]
        return [(cuts[i], cuts[i + 1]) for i in range(len(cuts) - 1)][::-1]

    def _forward_backward(self, grid_backward=True):
        s, lib, B = abi.stream(), abi.lib, self.B
        _ck(lib.ngp_step_begin(self.state.data_ptr(), self.beta1, self.beta2, s))
        if self.bin_bits:
            _ck(lib.ngp_bin_points(self.points.data_ptr(), self.sdfs.data_ptr(), B, 1.0, self.bin_bits, self.bin_ws.data_ptr(), self.bin_ws.numel(),
                                   self.pts.data_ptr(), self.tgt.data_ptr(), s))
        # encoder(x, bound=1): (x + 1) / 2, applied by the grid kernels as they load the points (ngp_call_opts.input_bound)
        _ck(lib.ngp_grid_encode_forward(self.pts.data_ptr(), self.table16.data_ptr(), self.offsets.data_ptr(), self.offsets_host.ctypes.data,
                                        self.enc.data_ptr(), B, 3, self.C, self.L, self.S, self.base_resolution, None, 0, 0, 0, abi.DTYPE_F16,
                                        abi.LAYOUT_BLC, s, abi.call_opts(input_bound=1.0)))
        _ck(lib.ngp_ffmlp_forward(self.enc.data_ptr(), self.mlp_w.data_ptr(), abi.DTYPE_F32, B, 32, 16, self.hid, self.nl, 0, 6, None, self.h.data_ptr(), s, None))
        _ck(lib.ngp_sdf_loss(self.h.data_ptr(), self.tgt.data_ptr(), B, self.state[_ST_SCALE:].data_ptr(), self.g_h.data_ptr(),
                             self.state[_ST_LOSS:].data_ptr(), s, None))
        _ck(lib.ngp_ffmlp_backward(self.g_h.data_ptr(), self.enc.data_ptr(), self.mlp_w.data_ptr(), abi.DTYPE_F32, None, B, 32, 16, self.hid, self.nl, 0, 6, 1,
                                   None, self.g_enc.data_ptr(), self.g32.data_ptr(), abi.DTYPE_F32, self.mlp_ws.data_ptr(), self.mlp_ws.numel(), s, None))
        if grid_backward:
            self._grid_backward()

    def _grid_backward(self, level_begin=0, level_end=0):
        s, lib, B = abi.stream(), abi.lib, self.B
        # batch left in its random order (bin_bits = 0): a few thousand coarse cells take every point -> those levels go through shared memory
        _ck(lib.ngp_grid_encode_backward(self.g_enc.data_ptr(), self.pts.data_ptr(), self.table16.data_ptr(), self.offsets.data_ptr(),
                                         self.offsets_host.ctypes.data, self.g16.data_ptr(), B, 3, self.C, self.L, self.S, self.base_resolution, None,
                                         None, 0, 0, 0, abi.DTYPE_F16, abi.LAYOUT_BLC, s,
                                         abi.call_opts(input_bound=1.0, flags=abi.OPT_GRID_BWD_SMEM_COARSE if self.coarse_smem else 0, level_begin=level_begin, level_end=level_end)))

    def _step_pipelined(self):
        """forward + MLP backward (one graph), then per level range: grid backward on the main stream | all-reduce of the
        finished range on the side stream; the optimizer (one graph) waits for the last range."""
        import torch.distributed as dist
        main = torch.cuda.current_stream()
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream()
            self._ev = [torch.cuda.Event() for _ in range(1 + self.L // 4)]
        reduce_ = self.world > 1 and dist.is_available() and dist.is_initialized()
        if self.use_graph and getattr(self, "_g_head", None) is None:
            self._forward_backward(grid_backward=False)  # first call eager: lazy initialisation happens outside the capture
            self._g_head = False
        elif self.use_graph:
            if self._g_head is False:
                torch.cuda.synchronize()
                self._g_head, self._g_opt = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._g_head):
                    self._forward_backward(grid_backward=False)
                with torch.cuda.graph(self._g_opt):
                    self._optimizer_step()
                # the two captures only recorded the launches: this step still has to run them
            self._g_head.replay()
        else:
            self._forward_backward(grid_backward=False)
        self._ev[0].record(main)
        off = self.offsets_host
        ranges = self._level_ranges()
        for i, (lb, le) in enumerate(ranges):
            self._grid_backward(lb, le)
            self._ev[i + 1].record(main)
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._ev[0])
            if reduce_:
                dist.all_reduce(self.g32.view(-1))
            for i, (lb, le) in enumerate(ranges):
                self._side.wait_event(self._ev[i + 1])
                if reduce_:
                    dist.all_reduce(self.g16[int(off[lb]):int(off[le])].view(-1))
        main.wait_stream(self._side)
        if self.use_graph and self._g_head not in (None, False):
            self._g_opt.replay()
        else:
            self._optimizer_step()

    def _optimizer_step(self):
        s, lib = abi.stream(), abi.lib
        _ck(lib.ngp_found_inf(self.g16.data_ptr(), self.n16, self.g32.data_ptr(), self.n32, self.state.data_ptr(), s))
        _ck(lib.ngp_adam_step(self.params.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.g16.data_ptr(), self.g32.data_ptr(),
                              self.table16.data_ptr(), self.n16, self.n32, self.state.data_ptr(), self.beta1, self.beta2, self.eps, float(self.world),
                              self.wd, s, None))
        _ck(lib.ngp_step_end(self.state.data_ptr(), s))

    def _step_eager(self):
        self._forward_backward()
        if self.world > 1:
            from .parallel import allreduce_gradients
            allreduce_gradients(self.g16, self.g32)
        self._optimizer_step()

    def train_step(self, points=None, sdfs=None):
        """points [B,3] in [-1,1], sdfs [B] (device or pinned host tensors); None keeps the current batch buffers."""
        if points is not None:  # host batches travel on a side stream while the previous step computes (nerf/feed.py)
            if getattr(self, "_feed", None) is None:
                from .feed import HostFeed
                self._feed = HostFeed(self.dev)
            self._feed.stage((self.points, self.sdfs), (points, sdfs))
        if self.grad_chunks > 1:
            self._step_pipelined()
        elif self.use_graph and self._graph is None:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                self._step_eager()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            self._g_fb, self._g_opt = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._g_fb):
                self._forward_backward()
            with torch.cuda.graph(self._g_opt):
                self._optimizer_step()
            self._graph = True
        elif self.use_graph:
            self._g_fb.replay()
            if self.world > 1:
                from .parallel import allreduce_gradients
                allreduce_gradients(self.g16, self.g32)
            self._g_opt.replay()
        else:
            self._step_eager()
        self.global_step += 1

    def loss(self):
        return float(self.state[_ST_LAST_LOSS].item())

    def loss_async(self):
        """Enqueue the device -> host copy of the last step's loss into pinned memory; returns a function that waits for THAT copy
        only and gives the value (collected one step later, the host keeps launching: see NGPTrainer.loss_async)."""
        ring = self.__dict__.setdefault("_loss_ring", {"i": 0, "buf": torch.zeros(8, dtype=torch.float32).pin_memory(),
                                                        "ev": [torch.cuda.Event() for _ in range(8)]})
        i = ring["i"] = (ring["i"] + 1) % 8
        ring["buf"][i:i + 1].copy_(self.state[_ST_LAST_LOSS:_ST_LAST_LOSS + 1], non_blocking=True)
        ev = ring["ev"][i]
        ev.record()

        def get():
            ev.synchronize()
            return float(ring["buf"][i])
        return get

    @torch.no_grad()
    def predict(self, points):
        n = points.shape[0]
        x01, enc, h = torch.empty(n, 3, device=self.dev), torch.empty(n, 32, dtype=torch.half, device=self.dev), torch.empty(n, 16, dtype=torch.half, device=self.dev)
        s = abi.stream()
        _ck(abi.lib.ngp_nerf_x01(points.contiguous().data_ptr(), 1.0, x01.data_ptr(), n, s, None))
        _ck(abi.lib.ngp_grid_encode_forward(x01.data_ptr(), self.table16.data_ptr(), self.offsets.data_ptr(), self.offsets_host.ctypes.data, enc.data_ptr(),
                                            n, 3, self.C, self.L, self.S, self.base_resolution, None, 0, 0, 0, abi.DTYPE_F16, abi.LAYOUT_BLC, s, None))
        _ck(abi.lib.ngp_ffmlp_inference(enc.data_ptr(), self.mlp_w.data_ptr(), abi.DTYPE_F32, n, 32, 16, self.hid, self.nl, 0, 6, None, h.data_ptr(), s, None))
        return h[:, 0].float()


# ---- synthetic data: analytic torus, sampled like sdf/provider.py:60-90 (surface, near-surface, uniform) -----------------
def torus_sdf(p, R=0.5, r=0.2):
    q = torch.stack([torch.sqrt(p[..., 0] ** 2 + p[..., 2] ** 2) - R, p[..., 1]], -1)
    return q.norm(dim=-1) - r


def sample_sdf_batch(n, device, seed=0, R=0.5, r=0.2):
    g = torch.Generator(device="cpu").manual_seed(seed)
    ns = n * 7 // 8
    u, v = torch.rand(ns, generator=g) * 2 * np.pi, torch.rand(ns, generator=g) * 2 * np.pi
    surf = torch.stack([(R + r * torch.cos(v)) * torch.cos(u), r * torch.sin(v), (R + r * torch.cos(v)) * torch.sin(u)], -1)
    surf[ns // 2:] += 0.01 * torch.randn(ns - ns // 2, 3, generator=g)  # perturbed surface points
    unif = torch.rand(n - ns, 3, generator=g) * 2 - 1
    p = torch.cat([surf, unif]).clamp(-1, 1).float().contiguous().to(device)
    return p, torus_sdf(p, R, r).contiguous()
