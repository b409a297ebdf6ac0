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
                 use_graph=True, num_layers=3, hidden_dim=64, log2_hashmap_size=19, grad_chunks=None, bin_bits=7, coarse_smem=None, overlap_optimizer=None, fused_loss=True):
        self.dev = dev = torch.device(device)
        self.B, self.world, self.use_graph = int(batch), world_size, use_graph
        self.beta1, self.beta2, self.eps, self.wd = 0.9, 0.99, 1e-15, weight_decay
        self.nl, self.hid = num_layers, hidden_dim
        # fused_loss: the net's forward and the loss in one launch (ngp_sdf_forward_loss; hidden 64 only): the prediction stays on the SM
        self.fused_loss = bool(fused_loss) and hidden_dim == 64
        # Table-gradient pipeline (default with world_size > 1: two ranges): the grid backward runs as `grad_chunks` level
        # ranges, finest first, and a side stream all-reduces the range that is done while the next one is still being
        # accumulated -- the step has no marcher to hide the collective behind (nerf/trainer.py).  Measured:
        # 2 ranges 1.2% faster than 1 on 2 and on 8 GPUs, 4 ranges slower; the first range's collective overlaps the second
        # range's backward only partly (the backward keeps every SM full).  grad_chunks = 1: one backward call, then the all-reduces.
        self.grad_chunks = int(grad_chunks) if grad_chunks else (2 if world_size > 1 else 1)
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
        self._graph, self._graphs, self._side, self._ev = False, {}, None, None
        self._pending, self._reduced, self._grads_ev, self._join_ev = False, False, None, None
        # Deferred optimizer: all-reduce + Adam of step n run on a side stream beside the binning of batch n+1.  Default: with
        # several ranks only -- on one GPU Adam (HBM-bound) and the binning kernels slow each other down by what the overlap
        # saves (measured: bin 75 -> 135 us, Adam 72 -> 99 us when run side by side).
        self.overlap_optimizer = world_size > 1 if overlap_optimizer is None else bool(overlap_optimizer)
        self.adam_blocks_per_sm = 0  # > 0: Adam's grid is capped so that it shares the SMs with the binning kernels (nerf/trainer.py)
        self.overlap_events = None   # a list: (side start, side end, main after bin, main after join) events per step, for the timeline
        self.global_step = 0

    def _level_ranges(self):
        """[(level_begin, level_end)] of the chunked backward, finest levels first (multiples of 4)."""
        n = max(1, min(self.grad_chunks, self.L // 4))
        per = -(-(self.L // 4) // n) * 4
        cuts = list(range(0, self.L, per)) + [self.L]
        return [(cuts[i], cuts[i + 1]) for i in range(len(cuts) - 1)][::-1]

    def _bin(self):
        """phase "bin": needs the batch only -- with a deferred optimizer it runs beside the previous step's all-reduce + Adam"""
        if self.bin_bits:
            _ck(abi.lib.ngp_bin_points(self.points.data_ptr(), self.sdfs.data_ptr(), self.B, 1.0, self.bin_bits, self.bin_ws.data_ptr(),
                                       self.bin_ws.numel(), self.pts.data_ptr(), self.tgt.data_ptr(), abi.stream()))

    def _forward_backward(self, grid_backward=True):
        self._bin()
        self._samples(grid_backward)

    def _samples(self, grid_backward=True):
        """phase "samples": forward, loss, backward (needs the parameters of the previous optimizer step)"""
        s, lib, B = abi.stream(), abi.lib, self.B
        _ck(lib.ngp_step_begin(self.state.data_ptr(), self.beta1, self.beta2, s))
        # encoder(x, bound=1): (x + 1) / 2, applied by the grid kernels as they load the points (ngp_call_opts.input_bound)
        _ck(lib.ngp_grid_encode_forward(self.pts.data_ptr(), self.table16.data_ptr(), self.offsets.data_ptr(), self.offsets_host.ctypes.data,
                                        self.enc.data_ptr(), B, 3, self.C, self.L, self.S, self.base_resolution, None, 0, 0, 0, abi.DTYPE_F16,
                                        abi.LAYOUT_BLC, s, abi.call_opts(input_bound=1.0)))
        if self.fused_loss:
            _ck(lib.ngp_sdf_forward_loss(self.enc.data_ptr(), self.mlp_w.data_ptr(), abi.DTYPE_F32, self.tgt.data_ptr(), B, self.nl,
                                         self.state[_ST_SCALE:].data_ptr(), self.g_h.data_ptr(), self.state[_ST_LOSS:].data_ptr(), s, None))
        else:
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
                                         abi.call_opts(input_bound=1.0, flags=(abi.OPT_GRID_BWD_SMEM_COARSE if self.coarse_smem else 0) | (abi.OPT_GRID_BWD_GROUP_RED if self.bin_bits else 0), level_begin=level_begin, level_end=level_end)))

    # ---- phases of a step; graph replays once captured -------------------------------------------------------------------
    def _run(self, what):
        if self.use_graph and self._graph:
            self._graphs[what].replay()
        else:
            {"bin": self._bin, "samples": self._samples, "head": lambda: self._samples(False), "opt": self._optimizer_step}[what]()

    def _dist(self):
        import torch.distributed as dist
        return dist if (self.world > 1 and dist.is_available() and dist.is_initialized()) else None

    def _backward(self):
        """Phase "samples".  grad_chunks > 1: the grid backward is issued per level range, finest first, and the side stream
        all-reduces a finished range while the main stream accumulates the next one; the optimizer then finds the gradients
        reduced (self._reduced) and only has to wait for the side stream."""
        if self.grad_chunks <= 1:
            self._run("samples")
            self._reduced = False
            return
        main, dist = torch.cuda.current_stream(), self._dist()
        if self._side is None:
            self._side = torch.cuda.Stream()
        if self._ev is None:
            self._ev = [torch.cuda.Event() for _ in range(1 + self.L // 4)]
        self._run("head")  # forward, loss, MLP backward
        self._ev[0].record(main)
        off, ranges = self.offsets_host, self._level_ranges()
        for i, (lb, le) in enumerate(ranges):
            self._grid_backward(lb, le)
            self._ev[i + 1].record(main)
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._ev[0])
            if dist:
                dist.all_reduce(self.g32.view(-1))
            for i, (lb, le) in enumerate(ranges):
                self._side.wait_event(self._ev[i + 1])
                if dist:
                    dist.all_reduce(self.g16[int(off[lb]):int(off[le])].view(-1))
        self._reduced = True

    def _apply(self):
        """all-reduce (unless the chunked backward has queued it already) + optimizer, on the current stream"""
        if self.world > 1 and not self._reduced:
            from .parallel import allreduce_gradients
            allreduce_gradients(self.g16, self.g32)
        self._run("opt")

    def flush(self):
        """Apply the deferred optimizer step, if any, on the current stream.  With several ranks this is a COLLECTIVE call
        (the all-reduce of the pending gradients): every rank has to make it."""
        if self._pending:
            if self._side is not None:
                torch.cuda.current_stream().wait_stream(self._side)
            self._apply()
            self._pending = False

    def _optimizer_step(self):
        s, lib = abi.stream(), abi.lib
        _ck(lib.ngp_found_inf(self.g16.data_ptr(), self.n16, self.g32.data_ptr(), self.n32, self.state.data_ptr(), s))
        _ck(lib.ngp_adam_step(self.params.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(), self.g16.data_ptr(), self.g32.data_ptr(),
                              self.table16.data_ptr(), self.n16, self.n32, self.state.data_ptr(), self.beta1, self.beta2, self.eps, float(self.world),
                              self.wd, s, abi.call_opts(grid_cap_blocks_per_sm=self.adam_blocks_per_sm) if self.adam_blocks_per_sm else None))
        _ck(lib.ngp_step_end(self.state.data_ptr(), s))

    def _step(self):
        main = torch.cuda.current_stream()
        if not self.overlap_optimizer:
            self.flush()
            self._run("bin")
            self._backward()
            self._pending = True
            self.flush()
            return
        # deferred optimizer: all-reduce + Adam of step n-1 on the side stream, beside the binning of batch n (which needs
        # the batch only); forward / backward of step n wait for both
        if self._pending:
            if self._side is None:
                self._side = torch.cuda.Stream()
            # the side stream waits for the gradients of the previous step -- for THAT event, not for everything queued on the
            # main stream since: a host batch staged by this call (nerf/feed.py) is waited for by the binning, not by the collective
            self._side.wait_event(self._grads_ev)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)] if self.overlap_events is not None else None
            if ev:
                ev[0].record(main)
            self._run("bin")
            if ev:
                ev[1].record(main)
            with torch.cuda.stream(self._side):
                if ev:
                    ev[2].record(self._side)
                self._apply()
                if ev:
                    ev[3].record(self._side)
            main.wait_stream(self._side)
            if self._join_ev is None:
                self._join_ev = torch.cuda.Event()
            self._join_ev.record(main)  # the next host batch starts its copy after the collective (nerf/feed.py `after`)
            if ev:
                ev[4].record(main)
                self.overlap_events.append(ev)
            self._pending = False
        else:
            self._run("bin")
        self._backward()
        if self._grads_ev is None:
            self._grads_ev = torch.cuda.Event()
        self._grads_ev.record(main)
        self._pending = True

    def _capture(self):
        """One eager step on a side stream (lazy kernel attributes, NCCL communicators), then the phases are captured.  The
        all-reduce stays outside the graphs."""
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            saved, self.use_graph = self.use_graph, False
            self._bin()
            self._backward()
            self._pending = True
            self.flush()
            self.use_graph = saved
        torch.cuda.current_stream().wait_stream(stream)
        torch.cuda.synchronize()
        phases = {"bin": self._bin, "samples": self._samples, "opt": self._optimizer_step}
        if self.grad_chunks > 1:
            phases["head"] = lambda: self._samples(False)
        self._graphs = {}
        for name, fn in phases.items():
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                fn()
            self._graphs[name] = g
        self._graph = True

    def train_step(self, points=None, sdfs=None):
        """points [B,3] in [-1,1], sdfs [B] (device or pinned host tensors); None keeps the current batch buffers.
        With overlap_optimizer (default) the optimizer step of this call is applied at the start of the next one, or by flush()."""
        if points is not None:  # host batches travel on a side stream while the previous step computes (nerf/feed.py)
            if getattr(self, "_feed", None) is None:
                from .feed import HostFeed
                self._feed = HostFeed(self.dev)
            self._feed.stage((self.points, self.sdfs), (points, sdfs), after=self._join_ev)
        if self.use_graph and not self._graph:
            self.flush()
            self._capture()  # its eager step is this call's step
        else:
            self._step()
        self.global_step += 1

    def loss(self):
        """Loss of the last step.  While its optimizer step is pending the value is still in the accumulator slot."""
        return float(self.state[_ST_LOSS if self._pending else _ST_LAST_LOSS].item())

    def loss_async(self):
        """Enqueue the device -> host copy of the last step's loss into pinned memory; returns a function that waits for THAT copy
        only and gives the value (collected one step later, the host keeps launching: see NGPTrainer.loss_async)."""
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

    @torch.no_grad()
    def predict(self, points):
        self.flush()
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
