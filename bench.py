"""bench.py -- the driver's measurement contract for the Instant-NGP hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One JSON line on stdout (rank 0).  Workloads (BASELINE.json `configs`):
    hashgrid    configs[0]  GridEncoder fwd+bwd, 2^18 random points, L=16 F=2 log2_hashmap=19, fp32
    nerf_train  configs[1]/[2]  NeRF training step (hash grid + cuda_ray + fp16), 4096 rays per GPU per step
    render      configs[3]  1920x1080 inference frame, row bands sharded over the GPUs
    sdf         configs[4]  hash-grid SDF + FFMLP training step, 2^20 points per GPU per step
A "step" is one pass of the hot path over one batch of synthetic input.  `value` is timed with inputs
resident in HBM; `e2e` is the same metric through the public plugin API with pinned HOST inputs, the
H2D copy and a D2H read of the result inside the timed region.  `--impl reference` times the CPU
restatement under oracle/ (the reference has no CPU implementation of this path) on
the host cores, bounded sample, rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "torch-ngp_b200"), os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402


def load_peaks():
    """Roofline denominators: MEASURED_PEAKS.json (driver-written) else the profiling recipe's fallback."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            p = json.load(f)
        return {"hbm_gbs": float(p["hbm_gbs"]), "tflops": float(p.get("bf16_tflops_sustained", p["bf16_tflops"])),
                "tflops_burst": float(p["bf16_tflops"]), "source": "measured"}
    except Exception:
        return {"hbm_gbs": 6650.0, "tflops": 1400.0, "tflops_burst": 1590.0, "source": "fallback"}


# Tensor-pipe utilisation of the FFMLP kernels is a profiler counter, not something a timed run can read: the committed
# `ncu --set full` capture (profiles/r2_ffmlp_ncu_summary.json) holds sm__pipe_tensor_cycles_active and
# sm__ops_path_tensor_src_fp16_dst_fp32 (the tcgen05-era counters; the hmma ones read n/a on UTCHMMA kernels).
def _tensor_pipe_note():
    try:
        with open(os.path.join(ROOT, "profiles", "r2_ffmlp_ncu_summary.json")) as f:
            cap = json.load(f)
        return {k: {"pipe_tensor_cycles_active_pct": v[0].get("tensor_pipe_active_pct"), "tensor_ops_pct_of_peak": v[0].get("tensor_ops_pct_of_peak"),
                    "us_under_ncu": v[0].get("us")} for k, v in cap["kernels"].items() if k.startswith("k_ffmlp")} | {"source": "profiles/r2_ffmlp_ncu_summary.json"}
    except Exception:
        return None


TENSOR_PIPE_NOTE = _tensor_pipe_note()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period_ms=100):
        self.index, self.period_ms, self.proc, self.lines = index, period_ms, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(n_gpus):
    """One process per GPU under torchrun; single process otherwise."""
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        if os.environ.get("NCCL_DEBUG", "").upper() in ("VERSION", "WARN"):
            os.environ.pop("NCCL_DEBUG")  # both levels print "NCCL version ..." on stdout; keep stdout to the one JSON line
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo", rank=rank, world_size=world)
    return rank, world, local


def barrier_sync(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def max_over_ranks(x, world):
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, world):
    if world == 1:
        return x
    import torch
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


# ==================================================================================================
# workload: GridEncoder fwd+bwd (BASELINE.json configs[0])
# ==================================================================================================
class HashGridWorkload:
    """2^18 points U[-1,1]^3 per GPU, L=16 F=2 log2_hashmap=19 desired_resolution=2048, fp32 table.
    Device step = forward + grad-table clear + backward through gridencoder._backend (C ABI).
    Algorithmic bytes: fwd 1,164 B/pt, bwd 1,164 B/pt + T*C*4 B table clear."""
    name = "hashgrid"
    l2_flush = True  # the 47 MB table and the 2^18-point buffers would sit in the 126 MB L2 from step to step
    metric = "hash-encode Mpts/s fwd+bwd"
    unit = "Mpts/s"
    dtype = "f32"
    scaling = "weak"

    def __init__(self, args):
        self.B = args.points or (1 << 18)
        self.L, self.C, self.D, self.H = 16, 2, 3, 16

    def config(self, world):
        return {"workload": "GridEncoder fwd+bwd, 2^18 random points in [-1,1]^3 per GPU, L=16 F=2 log2_hashmap=19 fp32 (configs[0])",
                "points_per_gpu": self.B, "table_entries": int(self.T), "l2": getattr(self, "l2_note", ""),
                "parallelism": f"dp{world} (independent points, no collective)"}

    def setup(self, dev, rank, world):
        import torch
        import gridencoder
        self.torch, self.dev = torch, dev
        torch.manual_seed(1000 + rank)
        self.enc = gridencoder.GridEncoder(desired_resolution=2048).to(dev)
        self.T = int(self.enc.offsets[-1].item())
        self.S = float(np.log2(self.enc.per_level_scale))
        self.x01 = torch.rand(self.B, 3, device=dev)
        self.grad = torch.randn(self.B, self.L * self.C, device=dev)
        self.out = torch.empty(self.B, self.L * self.C, device=dev)
        self.g_emb = torch.zeros_like(self.enc.embeddings)
        self.be = gridencoder._backend
        self.oh = self.be.host_offsets(self.enc.offsets)
        # e2e: pinned host inputs
        self.x_host = (self.x01 * 2 - 1).cpu().pin_memory()
        self.g_host = self.grad.cpu().pin_memory()
        self.x_dev = torch.empty_like(self.x01)
        self.units_per_step = self.B
        self.h2d_bytes = self.x_host.numel() * 4 + self.g_host.numel() * 4
        self.d2h_bytes = 4

    def step_device(self):
        import ngp_abi as abi
        e, emb = self.enc, self.enc.embeddings.data
        self.be.grid_encode_forward(self.x01, emb, e.offsets, self.out, self.B, 3, 2, self.L, self.S, self.H, None, 0, False, 0,
                                    layout=abi.LAYOUT_BLC, offsets_host=self.oh)
        self.g_emb.zero_()
        self.be.grid_encode_backward(self.grad, self.x01, emb, e.offsets, self.g_emb, self.B, 3, 2, self.L, self.S, self.H, None, None,
                                     0, False, 0, layout=abi.LAYOUT_BLC, offsets_host=self.oh)

    def step_e2e(self):
        """Public API: GridEncoder.forward + autograd backward; inputs and upstream gradient come from pinned host memory."""
        torch = self.torch
        self.x_dev.copy_(self.x_host, non_blocking=True)
        g = self.g_host.to(self.dev, non_blocking=True)
        self.enc.embeddings.grad = None
        y = self.enc(self.x_dev, bound=1)
        loss = (y * g).sum()
        loss.backward()
        return float(loss.item())  # D2H

    def launches_per_step(self):
        return None  # counted live

    def roofline(self, prof, peaks, steps):
        """Dominant kernel of the step (largest total device time among the grid kernels) against the HBM peak.
        Algorithmic bytes per launch (element size s=4): forward B*(12 + 8*L*C*s + L*C*s),
        backward B*(L*C*s + 12 + 8*L*C*s), times the share of the L levels that launch covers."""
        grid = {k: v for k, v in prof.items() if k.startswith("k_grid_")}
        if not grid:
            return None
        name = max(grid, key=lambda k: grid[k][1])
        cnt, ms = grid[name]
        share = 1.0
        if not (name.endswith("_lg4") or name.endswith("_v3")):  # the split variants: shared-memory levels vs L2 levels
            ns = self.n_smem_levels(prof)
            share = (ns if name.endswith("_smem") else self.L - ns) / self.L
        alg = self.B * (12 + 9 * self.L * self.C * 4) * share
        achieved = alg / (ms / cnt * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": name, "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": None, "peak_source": peaks["source"], "algorithmic_bytes_per_launch": alg, "us_per_launch": ms / cnt * 1e3}

    def n_smem_levels(self, prof):
        n = 0
        for l in range(self.L):
            if (int(self.oh[l + 1]) - int(self.oh[l])) * 8 <= 200 * 1024:
                n += 1
            else:
                break
        return n

    def step_algorithmic_bytes(self):
        s = 4
        return self.B * (4 * 3 + 8 * self.L * self.C * s + self.L * self.C * s) * 2 + self.T * self.C * s

    def reference_ext(self):
        """The reference's own CUDA extension (oracle/_ref/_gridencoder.so: gridencoder.cu compiled unmodified for sm_100a) on
        the SAME tensors, outside every timed region of this repository's path -- a reported baseline beside cpu_baseline."""
        try:
            from oracle import ref_ext
            ref = ref_ext.load("_gridencoder")
        except Exception as e:  # the checker is absent: say so, measure nothing
            return {"unavailable": f"{type(e).__name__}: {e}"[:200]}
        if ref is None:
            return {"unavailable": "oracle/_ref/_gridencoder.so not built"}
        torch = self.torch
        e, emb = self.enc, self.enc.embeddings.data
        out = torch.empty(self.L, self.B, self.C, device=self.dev)
        grad = self.grad.view(self.B, self.L, self.C).permute(1, 0, 2).contiguous()
        g_emb = torch.zeros_like(emb)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)
        offs = e.offsets.to(self.dev)

        def step():
            ref.grid_encode_forward(self.x01, emb, offs, out, self.B, 3, 2, self.L, self.S, self.H, None, 0, False, 0)
            g_emb.zero_()
            ref.grid_encode_backward(grad, self.x01, emb, offs, g_emb, self.B, 3, 2, self.L, self.S, self.H, None, None, 0, False, 0)
        for _ in range(3):
            step()
        ts = []
        for _ in range(10):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); step(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ms = float(np.median(ts))
        return {"value": self.B / (ms * 1e-3) / 1e6, "unit": self.unit, "ms_per_step": ms, "kind": "reference CUDA extension, unmodified, sm_100a build",
                "sample": "10 x fwd + table clear + bwd on the same 2^18 points and table, L2 flushed between steps, median"}

    # ---- CPU baseline: the oracle port (C + OpenMP), all host cores, bounded sample -------------
    def cpu_baseline(self, budget_s=15.0):
        from oracle import cpu_ref
        cpu_ref.build()
        offs, pls = cpu_ref.make_offsets(desired_resolution=2048)
        rng = np.random.default_rng(0)
        emb = ((rng.random((offs[-1], 2)) - 0.5) * 2e-4).astype(np.float32)
        Bs = 1 << 16
        x = rng.random((Bs, 3)).astype(np.float32)
        g = rng.standard_normal((16, Bs, 2)).astype(np.float32)
        t_total, n = 0.0, 0
        cpu_ref.grid_encode_forward(x, emb, offs, pls, 16)  # warm-up
        while t_total < budget_s and n < 64:
            t0 = time.perf_counter()
            cpu_ref.grid_encode_forward(x, emb, offs, pls, 16)
            cpu_ref.grid_encode_backward(g, x, emb, offs, pls, 16)
            t_total += time.perf_counter() - t0
            n += 1
        return {"value": Bs * n / t_total / 1e6, "unit": self.unit, "cores": os.cpu_count(), "kind": "port",
                "sample": f"{n} x fwd+bwd of 2^16 points (same table), oracle/ngp_oracle.c with OpenMP on all host cores"}


# ==================================================================================================
# workload: NeRF training step (BASELINE.json configs[1] at N=1, configs[2] at N>1)
# ==================================================================================================
class NerfTrainWorkload:
    """`main_nerf -O --ff` step on the synthetic gyroid scene: 4096 rays/batch on one GPU (configs[1]); with N>1 ranks
    2^18 rays/step global, data-parallel over rays, NCCL all-reduce of the table + MLP gradients (configs[2]).
    The model is pre-trained for `--pretrain` steps before the warm-up so that the occupancy grid (and with it the
    samples per ray) is that of a scene in training, not of the random initialisation."""
    name = "nerf_train"
    l2_flush = False  # a step streams ~470 MB (parameters + Adam moments 147 MB, fp16 table and gradient 49 MB, ~270 MB of sample buffers) > 126 MB L2
    metric = "train rays/s"
    unit = "rays/s"
    dtype = "f16"

    def __init__(self, args):
        self.args = args
        self.scaling = "weak" if args.rays else "strong"

    def setup(self, dev, rank, world):
        import torch
        import nerf
        from nerf import synthetic
        self.torch, self.dev, self.world = torch, dev, world
        self._overlap = not self.args.no_overlap
        torch.manual_seed(1000 + rank)  # ray jitter and density-grid probes; fp16 atomics still make runs differ in the last bits
        self.N = self.args.rays or (4096 if world == 1 else (1 << 18) // world)
        self.tr = tr = nerf.NGPTrainer(bound=1.0, num_rays=self.N, device=dev, world_size=world, seed=0, use_graph=not self.args.no_graph,
                                       overlap_optimizer=not self.args.no_overlap)
        tr.adam_blocks_per_sm = self.args.adam_blocks
        n_batches = 16
        views = max(16, n_batches * self.N // 4096)
        o, d, gt = synthetic.make_pool(views, n_batches * self.N // views, dev, seed=1000 + rank)
        tr.set_pool(o, d, gt)
        self.pool_host = tuple(t.cpu().pin_memory() for t in (o, d, gt))
        self.cursor = 0
        for _ in range(self.args.pretrain):
            tr.next_pool_batch()
            tr.train_step()
        torch.cuda.synchronize()
        self.units_per_step = self.N
        self.h2d_bytes = 3 * self.N * 3 * 4
        self.d2h_bytes = 4
        self.T = tr.T

    def config(self, world):
        tr = getattr(self, "tr", None)
        c = {"workload": ("main_nerf -O --ff training step (hash grid 16x2 T=2^19 + FFMLP 64x2 / 64x3 + cuda_ray marcher, fp16 + GradScaler + Adam), "
                          "synthetic 800x800 gyroid scene, " + ("4096 rays/batch, 1 GPU (configs[1])" if world == 1 and not self.args.rays else
                                                                "2^18 rays/step global, data-parallel (configs[2])" if not self.args.rays else
                                                                f"{self.args.rays} rays per GPU per step")),
             "parallelism": f"dp{world}" + (" (NCCL all-reduce of fp16 table + fp32 MLP gradients)" if world > 1 else ""),
             "l2": getattr(self, "l2_note", ""), "pretrain_steps": self.args.pretrain}
        if tr is not None:
            c.update({"rays_per_gpu": self.N, "global_rays": self.N * world, "samples_per_ray": round(tr.mean_count / self.N, 2),
                      "samples_per_step_per_gpu": tr.mean_count, "sample_capacity": tr.M, "cuda_graph": bool(tr.use_graph),
                      "optimizer_overlap": "optimizer(n-1) runs beside march(n) on a second stream, inside step n's timed region" if self._overlap else "none", "density_grid_update": f"every {tr.update_interval} steps, inside the timed region",
                      "last_loss": tr.loss()})
        return c

    def step_device(self):
        self.tr.next_pool_batch()
        self.tr.train_step()

    def step_e2e(self):
        """Public API: NGPTrainer.train_step(rays_o, rays_d, gt) with pinned host batches, loss read back."""
        o, d, g = self.pool_host
        if self.cursor + self.N > o.shape[0]:
            self.cursor = 0
        sl = slice(self.cursor, self.cursor + self.N)
        self.cursor += self.N
        self.tr.train_step(o[sl], d[sl], g[sl])
        # the loss of EVERY step is copied to the host inside the timed region (4 B, pinned); its value is collected one step
        # later (NGPTrainer.loss_async), so the host does not drain the launch queue once per step
        prev, self._loss_get = getattr(self, "_loss_get", None), self.tr.loss_async()
        return prev() if prev is not None else None

    def e2e_finish(self):
        """the last step's loss (waits for its copy): called inside the timed region, before the closing synchronize"""
        get, self._loss_get = getattr(self, "_loss_get", None), None
        return get() if get is not None else None

    def extras(self, steps, world):
        """N > 1: where the collective sits.  CUDA events on the two streams of `steps` more steps: the all-reduce and the optimizer
        run on the side stream beside march(n+1) on the main stream; `exposed` is what of them outlasts the march."""
        if world == 1 or not self._overlap:
            return {}
        torch, tr = self.torch, self.tr
        tr.overlap_events = []
        for _ in range(steps):
            self.step_device()
        torch.cuda.synchronize()
        ev, tr.overlap_events = tr.overlap_events, None
        if not ev:
            return {}
        ar = [e[0].elapsed_time(e[1]) * 1e3 for e in ev]
        opt = [e[1].elapsed_time(e[2]) * 1e3 for e in ev]
        exposed = [max(0.0, e[3].elapsed_time(e[2]) * 1e3) for e in ev]  # march end -> optimizer end (negative: fully hidden)
        loc = {"allreduce_us": float(np.mean(ar)), "optimizer_us": float(np.mean(opt)), "allreduce_plus_optimizer_exposed_us": float(np.mean(exposed))}
        return {"collective": {k: max_over_ranks(v, world) for k, v in loc.items()} |
                {"note": "CUDA events on the side stream (all-reduce, optimizer) and the main stream (march of the next step), max over ranks of the per-rank means; "
                         "exposed = side-stream end minus march end, clamped at 0"}}

    def single_gpu_same_global_batch(self, rank, world, steps, warmup):
        """SURVEY.md 8(d) config 3's scaling denominator: rank 0 ALONE runs the global batch (2^18 rays/step) on one GPU with the
        model state (parameters, occupancy grid) the group has just trained, while the other ranks wait at a barrier."""
        import torch
        import nerf
        from nerf import synthetic
        out = None
        # collective part first, on EVERY rank: a deferred optimizer step of the group trainer carries an all-reduce (flush), and
        # state_dict() would otherwise run it on rank 0 alone -- against the other ranks' barrier
        self.tr.flush()
        barrier_sync(world)
        if rank == 0:
            Ng = self.N * world
            spr = max(8.0, self.tr.mean_count / self.N) if self.tr.mean_count > 0 else 256.0
            tr = nerf.NGPTrainer(bound=1.0, num_rays=Ng, device=self.dev, world_size=1, seed=0, use_graph=not self.args.no_graph,
                                 overlap_optimizer=not self.args.no_overlap, init_samples_per_ray=int(spr * 1.3) + 8)
            tr.adam_blocks_per_sm = self.args.adam_blocks
            tr.load_state_dict(self.tr.state_dict())
            tr.exp_avg.copy_(self.tr.exp_avg); tr.exp_avg_sq.copy_(self.tr.exp_avg_sq); tr.state.copy_(self.tr.state)
            tr.mean_sum.copy_(self.tr.mean_sum)
            tr.iter_density, tr.update_interval = self.tr.iter_density, 0  # the group's grid as it is: identical occupancy, so identical samples per ray
            views = max(16, 4 * Ng // 4096)
            o, d, gt = synthetic.make_pool(views, 4 * Ng // views, self.dev, seed=1000)
            tr.set_pool(o, d, gt)
            for _ in range(max(3, warmup)):
                tr.next_pool_batch(); tr.train_step()
            torch.cuda.synchronize()
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for a, b in ev:
                a.record(); tr.next_pool_batch(); tr.train_step(); b.record()
            torch.cuda.synchronize()
            ms = sum(a.elapsed_time(b) for a, b in ev) / steps
            cnt = int(tr.step_counter[0, 0].item())
            out = {"value": Ng / (ms * 1e-3), "unit": self.unit, "ms_per_step": ms, "rays_per_step": Ng, "n_gpus": 1, "steps": steps,
                   "samples_per_step": cnt, "note": "rank 0 alone, same parameters and occupancy grid as the group run, density-grid update off during the timed steps"}
            del tr
            torch.cuda.empty_cache()
        barrier_sync(world)
        return out

    def profile_mode(self, on):
        """per-launch CUDA events cannot be recorded inside a graph replay: the profile pass runs the same launches eagerly."""
        self._saved = getattr(self, "_saved", self.tr.use_graph)
        self.tr.use_graph = False if on else self._saved
        if hasattr(self.tr, "flush"):
            if on:  # per-kernel durations are taken with the kernels one after the other
                self.tr.flush()
            self.tr.overlap_optimizer = False if on else self._overlap

    def launches_per_step(self):
        """Kernels of this library in one step (graph replays do not pass through the host-side counter)."""
        import ngp_abi as abi
        saved, self.tr.use_graph = self.tr.use_graph, False
        upd, self.tr.update_interval = self.tr.update_interval, 0
        n0 = abi.launch_count()
        self.step_device()
        n = abi.launch_count() - n0
        self.tr.use_graph, self.tr.update_interval = saved, upd
        self.tr.global_step -= 1  # keep the density-grid cadence of the timed region
        return n

    def roofline(self, prof, peaks, steps):
        """Dominant kernel (largest share of the step) against the roofline that bounds it.  ALGORITHMIC work per launch
        (SURVEY.md 8d), S = samples actually marched per step (not the padded capacity), s = 2 B (fp16 table / activations):
          k_grid_fwd / k_grid_bwd  S * (12 + 8*L*C*s + L*C*s) bytes
          k_adam                   n16 * (3*4 read + 3*4 write + 2 grad + 2 shadow) + n32 * 32 bytes
          k_found_inf              n16 * 2 + n32 * 4 bytes
          k_ffmlp_fwd              2 * W * S flop
          k_ffmlp_bwd              4 * W * S flop (dL/dx and dL/dW; + nothing for the forward the kernel RECOMPUTES per tile --
                                   that work is reported separately as frac_incl_recompute = 6 * W * S)
        W = weights of the net, averaged over the sigma and colour nets (one launch each per step)."""
        tr = self.tr
        S = max(1, tr.mean_count)
        w_avg = (tr.n_sig + tr.n_col) / 2
        hb = S * (12 + 9 * tr.L * tr.C * 2)
        table = {"k_grid_fwd": ("hbm", hb), "k_grid_bwd": ("hbm", hb), "k_adam": ("hbm", tr.n16 * 28 + tr.n32 * 32),
                 "k_found_inf": ("hbm", tr.n16 * 2 + tr.n32 * 4),
                 "k_ffmlp_fwd": ("tensor", 2 * w_avg * S), "k_ffmlp_bwd": ("tensor", 4 * w_avg * S)}
        cand = {k: v for k, v in prof.items() if any(k.startswith(p) for p in table)}
        if not cand:
            return None
        name = max(cand, key=lambda k: cand[k][1])
        cnt, ms = cand[name]
        bound, work = next(v for p, v in table.items() if name.startswith(p))
        sec = ms / cnt * 1e-3
        total_ms = sum(v[1] for v in prof.values())
        if bound == "hbm":
            ach, peak, unit = work / sec / 1e9, peaks["hbm_gbs"], "GB/s"
        else:
            ach, peak, unit = work / sec / 1e12, peaks["tflops"], "TFLOP/s"
        r = {"bound": bound, "kernel": name, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak, "traffic": self._traffic(name, S),
             "peak_source": peaks["source"], "algorithmic_work_per_launch": work, "us_per_launch": sec * 1e6,
             "share_of_step_kernel_time": ms / total_ms if total_ms else None, "samples_per_step": S}
        if name.startswith("k_ffmlp_bwd"):
            r["frac_incl_recompute"] = ach * 1.5 / peak  # 6 W S: what the tensor pipe executes (forward recomputed on the SM)
        r["tensor_pipe"] = TENSOR_PIPE_NOTE
        return r

    def step_roofline(self, ms, peaks):
        """Whole training step against the HBM roofline, SURVEY.md 8(d) config 2: (N * S_bar * 1,656 B + N * 120 B) / t_step."""
        S = max(1, self.tr.mean_count)
        a = S * 1656 + self.N * 120
        ach = a / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"],
                "algorithmic_bytes_per_step": a, "formula": "samples * 1656 B + rays * 120 B (SURVEY.md 8d, config 2)"}

    def _traffic(self, name, S):
        """DRAM bytes (read + write) per launch of that kernel from the committed `ncu --set full` capture
        (profiles/r2_nerf_train_ncu_summary.json, else the round-1 one), scaled per sample to this run's sample count; None if absent."""
        try:
            path = os.path.join(ROOT, "profiles", "r2_nerf_train_ncu_summary.json")
            if not os.path.exists(path):
                path = os.path.join(ROOT, "profiles", "r1_nerf_train_ncu_summary.json")
            with open(path) as f:
                cap = json.load(f)
            runs = next(v for k, v in cap["kernels"].items() if name.startswith(k))
            per = sum(r["dram_read_bytes"] + r["dram_write_bytes"] for r in runs) / len(runs)
            return per if name.startswith(("k_adam", "k_found_inf")) else per * S / cap["samples"]
        except Exception:
            return None

    # ---- CPU baseline: the oracle's restatement of the same step, all host cores, bounded sample ----
    def cpu_baseline(self, budget_s=15.0):
        from oracle import cpu_ref, nerf_step_ref as R
        import scenes
        cpu_ref.build()
        offs, pls = cpu_ref.make_offsets(desired_resolution=2048)
        rng = np.random.default_rng(0)
        table = ((rng.random((offs[-1], 2)) * 2 - 1) * 1e-4).astype(np.float32)
        k = np.sqrt(3 / 64)
        st = R.RefState(table, ((rng.random(64 * (32 + 64 + 16)) * 2 - 1) * k).astype(np.float32),
                        ((rng.random(64 * (32 + 128 + 16)) * 2 - 1) * k).astype(np.float32), offs, pls, 16)
        H, n_rays = 128, 512
        bits = scenes.bitfield_from_grid(scenes.density_grid(H=H), 0.01)
        batches = []
        for i in range(4):
            ro, rd = scenes.camera_rays(n_rays, seed=i)
            batches.append((ro, rd, rng.random((n_rays, 3)).astype(np.float32)))
        R.train_step(st, *batches[0], bits, 1, H, n_rays * 128)  # warm-up
        t_total, n, samples = 0.0, 0, 0
        while t_total < budget_s and n < 64:
            t0 = time.perf_counter()
            _, ns = R.train_step(st, *batches[n % 4], bits, 1, H, n_rays * 128)
            t_total += time.perf_counter() - t0
            n += 1
            samples += ns
        return {"value": n_rays * n / t_total, "unit": self.unit, "cores": os.cpu_count(), "kind": "port",
                "sample": f"{n} training steps of {n_rays} rays ({samples // max(1, n)} samples/step, analytic occupancy grid, full-size table and MLPs), "
                          "oracle/nerf_step_ref.py: C+OpenMP hash grid / marcher / compositor, numpy MLPs and Adam"}


# ==================================================================================================
# workload: inference rendering (BASELINE.json configs[3])
# ==================================================================================================
class RenderWorkload:
    """1920x1080 frames of the synthetic scene from a 128^3 density bitfield: march_rays / composite_rays loop with the
    network_ff model (briefly pre-trained so that the occupancy grid is that of a scene).  Each rank renders one
    share of the rows (nerf/parallel.py::row_stripes), no collective; a step is one frame (one share per rank)."""
    name = "render"
    l2_flush = False  # a frame streams ~100 GB through the sample buffers (~160 M samples x ~600 B)
    metric = "render Mrays/s"
    unit = "Mrays/s"
    dtype = "f16"
    scaling = "strong"

    def __init__(self, args):
        self.args = args
        self.W, self.Hh = (1920, 1080) if not args.points else (args.points, args.points * 9 // 16)

    def setup(self, dev, rank, world):
        import torch
        import nerf
        from nerf import synthetic
        from nerf.parallel import row_stripes
        self.torch, self.dev = torch, dev
        torch.manual_seed(7)
        self.tr = tr = nerf.NGPTrainer(bound=1.0, num_rays=4096, device=dev, seed=0, use_graph=True)
        o, d, gt = synthetic.make_pool(16, 4096, dev, seed=1)
        tr.set_pool(o, d, gt)
        for _ in range(self.args.pretrain):
            tr.next_pool_batch()
            tr.train_step()
        self.rows = row_stripes(self.Hh, rank, world, stripe=8)  # interleaved 8-row stripes: every rank gets its share of the object
        self.frames = [synthetic.frame_rays(self.W, self.Hh, dev, pose_index=i * 7, rows=self.rows.numpy()) for i in range(4)]
        self.frames_host = [(o.cpu().pin_memory(), d.cpu().pin_memory()) for o, d in self.frames]
        self.i = 0
        self.units_per_step = int(self.rows.numel()) * self.W
        self.h2d_bytes = self.units_per_step * 6 * 4
        self.d2h_bytes = self.units_per_step * 4 * 4
        self.T = tr.T
        self.out_host = torch.empty(self.units_per_step, 3).pin_memory()
        self.depth_host = torch.empty(self.units_per_step).pin_memory()
        self.work_sum, self.frames_done = torch.zeros(2, dtype=torch.int64, device=dev), 0

    def config(self, world):
        c = {"workload": f"inference render {self.W}x{self.Hh} frames, march_rays/composite_rays from a 128^3 bitfield + network_ff shading (configs[3])",
             "parallelism": f"rows dealt to {world} rank(s) in interleaved 8-row stripes, no collective", "l2": getattr(self, "l2_note", ""),
             "pretrain_steps": self.args.pretrain}
        if hasattr(self, "tr"):
            c.update({"rows_per_gpu": int(self.rows.numel()), "loop_iterations": getattr(self.tr, "last_render_iters", None)})
        return c

    def step_device(self):
        o, d = self.frames[self.i % 4]
        self.i += 1
        self.tr.render(o, d, use_graph=getattr(self, "_graph_ok", True))
        self.work_sum += self.tr.last_render_work
        self.frames_done += 1

    def profile_mode(self, on):
        """per-launch CUDA events cannot be recorded inside a graph capture / replay: the profile pass runs the loop eagerly"""
        self._graph_ok = not on

    def step_e2e(self):
        o, d = self.frames_host[self.i % 4]
        self.i += 1
        img, depth = self.tr.render(o.to(self.dev, non_blocking=True), d.to(self.dev, non_blocking=True))
        self.out_host.copy_(img, non_blocking=True)
        self.depth_host.copy_(depth, non_blocking=True)
        self.torch.cuda.synchronize()
        return float(self.out_host[0, 0])

    # algorithmic bytes: per emitted sample 32 B march write + 24 B composite read + 588 B hash encode
    # (fp16 table: 12 + 8*L*C*2 + L*C*2) + 192 B MLP in/out; per ray and loop iteration 44 B read + 28 B state read-modify-write
    SAMPLE_BYTES = {"k_march_rays": 32, "k_composite_rays": 24, "k_grid_fwd": 588, "k_ffmlp_fwd": 96}
    RAY_ITER_BYTES = {"k_march_rays": 44, "k_composite_rays": 28}

    def _frame_work(self):
        """(sample rows shaded, ray-iterations marched) per frame: device-side totals the render loop keeps, averaged over every frame so far"""
        w = self.work_sum.cpu().double() / max(1, self.frames_done)
        return float(w[0]), float(w[1])

    def roofline(self, prof, peaks, steps):
        cand = {k: v for k, v in prof.items() if k.startswith(tuple(self.SAMPLE_BYTES))}
        if not cand:
            return None
        name = max(cand, key=lambda k: cand[k][1])
        key = next(k for k in self.SAMPLE_BYTES if name.startswith(k))
        cnt, ms = cand[name]
        total_ms = sum(v[1] for v in prof.values())
        rows, ray_iters = self._frame_work()
        alg = rows * self.SAMPLE_BYTES[key] + ray_iters * self.RAY_ITER_BYTES.get(key, 0)  # per frame, all launches of that kernel
        ach = alg / (ms / steps * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": name, "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"], "traffic": None,
                "peak_source": peaks["source"], "us_per_launch": ms / cnt * 1e3, "launches_per_step": cnt / steps,
                "algorithmic_bytes_per_step": alg, "samples_per_frame": rows, "ray_iterations_per_frame": ray_iters,
                "share_of_step_kernel_time": ms / total_ms if total_ms else None,
                "note": "all launches of the kernel in one frame together: achieved = algorithmic bytes of the frame's samples / their summed duration"}

    def step_roofline(self, ms, peaks):
        rows, ray_iters = self._frame_work()
        a = rows * sum(self.SAMPLE_BYTES.values()) + rows * 96 + ray_iters * sum(self.RAY_ITER_BYTES.values())  # two MLP launches per sample: 2 x 96 B
        ach = a / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": ach, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": ach / peaks["hbm_gbs"], "algorithmic_bytes_per_step": a,
                "formula": "samples * (32 + 24 + 588 + 192) B + ray-iterations * (44 + 28) B (SURVEY.md 8d, config 4b)"}

    def cpu_baseline(self, budget_s=15.0):
        from oracle import cpu_ref, nerf_step_ref as R
        import scenes
        cpu_ref.build()
        offs, pls = cpu_ref.make_offsets(desired_resolution=2048)
        rng = np.random.default_rng(0)
        k = np.sqrt(3 / 64)
        st = R.RefState(((rng.random((offs[-1], 2)) * 2 - 1) * 1e-4).astype(np.float32), ((rng.random(64 * 112) * 2 - 1) * k).astype(np.float32),
                        ((rng.random(64 * 176) * 2 - 1) * k).astype(np.float32), offs, pls, 16)
        bits = scenes.bitfield_from_grid(scenes.density_grid(H=128), 0.01)
        W, Hh = 96, 54
        ro, rd = scenes.frame_rays(W, Hh, pose_index=3)
        t_total, n = 0.0, 0
        while t_total < budget_s and n < 16:
            t0 = time.perf_counter()
            R.render(st, ro, rd, bits, 1, 128)
            t_total += time.perf_counter() - t0
            n += 1
        return {"value": W * Hh * n / t_total / 1e6, "unit": self.unit, "cores": os.cpu_count(), "kind": "port",
                "sample": f"{n} frames of {W}x{Hh} rays (1/400 of a 1920x1080 frame), oracle/nerf_step_ref.py::render on the analytic occupancy grid"}


# ==================================================================================================
# workload: hash-grid SDF training step with the FFMLP tensor-core path (BASELINE.json configs[4])
# ==================================================================================================
class SDFWorkload:
    """main_sdf --fp16 --ff on an analytic torus: 2^20 points per GPU per step (surface / near-surface / uniform samples),
    hash grid 16x2 T=2^19 -> FFMLP 32->64x3->1, mape loss, GradScaler + Adam; data-parallel over points with N>1."""
    name = "sdf"
    l2_flush = False  # 2^20 points x ~330 B of sample buffers + 196 MB of parameters, moments and gradients > 126 MB L2
    metric = "sdf train Mpts/s"
    unit = "Mpts/s"
    dtype = "f16"
    scaling = "weak"

    def __init__(self, args):
        self.args = args
        self.B = args.points or (1 << 20)

    def setup(self, dev, rank, world):
        import torch
        from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
        self.torch, self.dev = torch, dev
        self.tr = SDFTrainer(batch=self.B, device=dev, world_size=world, use_graph=not self.args.no_graph,
                             grad_chunks=getattr(self.args, "sdf_chunks", 0) or None,
                             overlap_optimizer=False if self.args.no_overlap else None,
                             **({} if getattr(self.args, "sdf_bin_bits", -1) < 0 else {"bin_bits": self.args.sdf_bin_bits}))
        self.batches = [sample_sdf_batch(self.B, dev, seed=100 * rank + i) for i in range(4)]
        self.batches_host = [(p.cpu().pin_memory(), y.cpu().pin_memory()) for p, y in self.batches]
        self.i = 0
        for _ in range(8):
            self.step_device()
        self.units_per_step = self.B
        self.h2d_bytes, self.d2h_bytes = self.B * 16, 4
        self.T = self.tr.T

    def config(self, world):
        c = {"workload": "main_sdf --fp16 --ff training step, analytic torus SDF, 2^20 points/batch per GPU (configs[4])",
             "parallelism": f"dp{world}" + (" (NCCL all-reduce of table + MLP gradients)" if world > 1 else ""),
             "l2": getattr(self, "l2_note", ""), "points_per_gpu": self.B}
        if hasattr(self, "tr"):
            c["last_loss"] = self.tr.loss()
            c["table_gradient_level_ranges"] = self.tr._level_ranges()
            c["batch_binned_by_morton_bits"] = self.tr.bin_bits
            c["optimizer_deferred_beside_next_binning"] = self.tr.overlap_optimizer
        return c

    def step_device(self):
        p, y = self.batches[self.i % 4]
        self.i += 1
        self.tr.train_step(p, y)

    def step_e2e(self):
        p, y = self.batches_host[self.i % 4]
        self.i += 1
        self.tr.train_step(p, y)  # 16 MB of pinned host input per step, staged on a side stream behind the previous step
        prev, self._loss_get = getattr(self, "_loss_get", None), self.tr.loss_async()  # every step's loss is copied; collected one step later
        return prev() if prev is not None else None

    def e2e_finish(self):
        get, self._loss_get = getattr(self, "_loss_get", None), None
        return get() if get is not None else None

    def profile_mode(self, on):
        self._saved = getattr(self, "_saved", self.tr.use_graph)
        self.tr.use_graph = False if on else self._saved
        if hasattr(self.tr, "flush"):  # a trainer with a deferred optimizer: per-kernel durations are taken one after the other
            if on:
                self.tr.flush()
                self._saved_overlap = self.tr.overlap_optimizer
            self.tr.overlap_optimizer = False if on else getattr(self, "_saved_overlap", False)

    def launches_per_step(self):
        import ngp_abi as abi
        saved, self.tr.use_graph = self.tr.use_graph, False
        n0 = abi.launch_count()
        self.step_device()
        n = abi.launch_count() - n0
        self.tr.use_graph = saved
        return n

    def extras(self, steps, world):
        """N > 1: CUDA-event timeline of the side stream (all-reduce + optimizer of step n) against the binning of batch n+1 on
        the main stream -- every rank runs the steps (collective), rank 0 reports its own events."""
        if world == 1 or not self.tr.overlap_optimizer:
            return {}
        self.tr.overlap_events = []
        for _ in range(12):
            self.step_device()
        self.torch.cuda.synchronize()
        ev, self.tr.overlap_events = self.tr.overlap_events[2:], None
        f = lambda i, j: sum(e[i].elapsed_time(e[j]) for e in ev) / len(ev) * 1e3
        return {"collective": {"allreduce_plus_optimizer_us": f(2, 3), "binning_beside_it_us": f(0, 1), "joined_us": f(0, 4),
                               "exposed_us": f(0, 4) - f(0, 1),
                               "note": "side stream: NCCL all-reduce of the fp16 table gradient + fp32 MLP gradients, found_inf, Adam; main stream: "
                                       "ngp_bin_points of the next batch; exposed = joined - binning"}}

    def roofline(self, prof, peaks, steps):
        tr, B = self.tr, self.B
        hb = B * (12 + 9 * tr.L * tr.C * 2)
        table = {"k_grid_fwd": ("hbm", hb), "k_grid_bwd": ("hbm", hb), "k_adam": ("hbm", tr.n16 * 28 + tr.n32 * 32),
                 "k_ffmlp_fwd": ("tensor", 2 * tr.n32 * B), "k_ffmlp_bwd": ("tensor", 4 * tr.n32 * B)}  # algorithmic: recompute not counted
        cand = {k: v for k, v in prof.items() if any(k.startswith(p) for p in table)}
        if not cand:
            return None
        name = max(cand, key=lambda k: cand[k][1])
        cnt, ms = cand[name]
        bound, work = next(v for p, v in table.items() if name.startswith(p))
        sec = ms / cnt * 1e-3
        total_ms = sum(v[1] for v in prof.values())
        ach, peak, unit = (work / sec / 1e9, peaks["hbm_gbs"], "GB/s") if bound == "hbm" else (work / sec / 1e12, peaks["tflops"], "TFLOP/s")
        return {"bound": bound, "kernel": name, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak, "traffic": None,
                "peak_source": peaks["source"], "algorithmic_work_per_launch": work, "us_per_launch": sec * 1e6,
                "share_of_step_kernel_time": ms / total_ms if total_ms else None}

    def cpu_baseline(self, budget_s=15.0):
        from oracle import cpu_ref, nerf_step_ref as R
        cpu_ref.build()
        offs, pls = cpu_ref.make_offsets(desired_resolution=2048)
        rng = np.random.default_rng(0)
        st = R.SDFRefState(((rng.random((offs[-1], 2)) * 2 - 1) * 1e-4).astype(np.float32),
                           ((rng.random(64 * (32 + 128 + 16)) * 2 - 1) * np.sqrt(3 / 64)).astype(np.float32), offs, pls, 16)
        Bs = 1 << 15
        p = (rng.random((Bs, 3)) * 2 - 1).astype(np.float32)
        y = (np.linalg.norm(np.stack([np.hypot(p[:, 0], p[:, 2]) - 0.5, p[:, 1]], -1), axis=-1) - 0.2).astype(np.float32)
        loss, grads = R.sdf_forward_backward(st, p, y)
        t_total, n = 0.0, 0
        while t_total < budget_s and n < 32:
            t0 = time.perf_counter()
            loss, grads = R.sdf_forward_backward(st, p, y)
            R.sdf_adam_step(st, grads)
            t_total += time.perf_counter() - t0
            n += 1
        return {"value": Bs * n / t_total / 1e6, "unit": self.unit, "cores": os.cpu_count(), "kind": "port",
                "sample": f"{n} training steps of 2^15 points (full-size table and MLP), oracle/nerf_step_ref.py::sdf_forward_backward + sdf_adam_step"}


WORKLOADS = {"hashgrid": HashGridWorkload, "nerf_train": NerfTrainWorkload, "render": RenderWorkload, "sdf": SDFWorkload}
DEFAULT_WORKLOAD = "nerf_train"


# ==================================================================================================
# driver
# ==================================================================================================
def timed_region(wl, steps, world, flush):
    """EXACTLY `steps` steps, barrier + synchronize on both sides, per-step CUDA events on the launching
    (current) stream; the L2 flush sits between the per-step event pairs.  -> mean ms per step (this rank)."""
    import torch
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    barrier_sync(world)
    for a, b in ev:
        if flush is not None:
            flush.zero_()
        a.record()
        wl.step_device()
        b.record()
    barrier_sync(world)
    wl.last_step_ms = [a.elapsed_time(b) for a, b in ev]
    return sum(wl.last_step_ms) / steps


def e2e_region(wl, steps, world):
    """Same metric through the public API with pinned host inputs; wall clock around H2D + step + D2H."""
    barrier_sync(world)
    t0 = time.perf_counter()
    for _ in range(steps):
        wl.step_e2e()
    if hasattr(wl, "e2e_finish"):
        wl.e2e_finish()
    barrier_sync(world)
    return (time.perf_counter() - t0) * 1e3 / steps


def run_reference(wl, args, rank, world):
    """`--impl reference`: the CPU restatement of the path on the host cores (rank 0 only)."""
    if rank != 0:
        return
    per = max(2.0, min(20.0, 60.0 / max(1, args.steps + args.warmup)))
    for _ in range(min(args.warmup, 1)):
        wl.cpu_baseline(budget_s=1.0)
    vals = [wl.cpu_baseline(budget_s=per) for _ in range(max(1, min(args.steps, 4)))]
    v = float(np.median([x["value"] for x in vals]))
    cb = dict(vals[0]); cb["value"] = v
    if not hasattr(wl, "T"):
        wl.T = 6119864
    line = {"impl": "reference", "metric": wl.metric, "value": v, "unit": wl.unit, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": None, "higher_is_better": True, "scaling": wl.scaling, "vs_baseline": None,
            "dtype": wl.dtype, "data": "synthetic",
            "config": dict(wl.config(world), reference_arm_sample=cb.get("sample"),
                           reference_arm_note="CPU restatement timed on a BOUNDED sample of the workload (see reference_arm_sample); the value is the "
                                              "sample's own throughput, i.e. an extrapolation to the full batch assuming linear cost in rays / points"),
            "cpu_baseline": cb, "e2e": {"value": v, "unit": wl.unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def measure(wl, args, rank, world, dev, peaks, steps, warmup, cpu_baseline=True, cpu_budget_s=15.0):
    """One workload, the whole contract: warm-up, EXACTLY `steps` timed steps (device events, barrier + synchronize on both
    sides, max over ranks), the per-kernel live profile of the same region, the end-to-end leg through the plugin API with
    pinned host buffers, and (rank 0, one GPU) the CPU baseline.  -> the JSON line as a dict (rank 0), None elsewhere."""
    import torch
    import ngp_abi as abi
    wl.setup(dev, rank, world)
    do_flush = (getattr(wl, "l2_flush", True) or args.flush) and not args.no_flush
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if do_flush else None
    wl.l2_note = ("flushed between steps (256 MiB memset outside the per-step events)" if do_flush else
                  "not flushed: the data one step streams (see workload) is several times the 126 MB L2")
    for _ in range(warmup):
        wl.step_device()
    torch.cuda.synchronize()

    sampler = ClockSampler(index=dev.index or 0) if rank == 0 else None
    if sampler:
        sampler.start()
    n0 = abi.launch_count()
    ms_local = timed_region(wl, steps, world, flush)
    per_step = sorted(wl.last_step_ms)
    step_ms = {"median": per_step[len(per_step) // 2], "min": per_step[0], "max": per_step[-1],
               "note": "per-step device durations of the timed region on rank 0 (ms_per_step is their mean, max over ranks); a step that "
                       "contains host work with read-backs (nerf_train: the density-grid update every 16th step) shows up as the max"}
    launches = (abi.launch_count() - n0) / steps
    if hasattr(wl, "launches_per_step") and wl.launches_per_step.__func__ is not HashGridWorkload.launches_per_step:
        launches = wl.launches_per_step()
    clocks = sampler.stop() if sampler else None
    ms = max_over_ranks(ms_local, world)
    units = sum_over_ranks(float(wl.units_per_step), world)

    # per-kernel durations, live, same region re-run with the library's per-launch CUDA events enabled
    if hasattr(wl, "profile_mode"):
        wl.profile_mode(True)
    abi.set_option("profile_events", 1)
    timed_region(wl, steps, world, flush)
    prof = abi.profile_collect()
    abi.set_option("profile_events", 0)
    if hasattr(wl, "profile_mode"):
        wl.profile_mode(False)
    roof = wl.roofline(prof, peaks, steps)

    # end to end through the plugin API, host buffers
    for _ in range(2):
        wl.step_e2e()
    if hasattr(wl, "e2e_finish"):
        wl.e2e_finish()
    ms_e2e = max_over_ranks(e2e_region(wl, steps, world), world)
    extra = wl.extras(steps, world) if hasattr(wl, "extras") else {}

    if rank != 0:
        return None
    scale = 1e-6 if wl.unit.startswith("M") else 1.0
    line = {"metric": wl.metric, "value": units / (ms * 1e-3) * scale, "unit": wl.unit, "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": wl.scaling, "vs_baseline": None,
            "dtype": wl.dtype, "data": "synthetic", "config": wl.config(world),
            "e2e": {"value": units / (ms_e2e * 1e-3) * scale, "unit": wl.unit, "h2d_bytes_per_step": wl.h2d_bytes,
                    "d2h_bytes_per_step": wl.d2h_bytes, "ms_per_step": ms_e2e},
            "gpu_launches": launches, "clocks": clocks, "step_ms": step_ms, "roofline": roof,
            "kernels": {k: {"launches_per_step": c / steps, "us_per_launch": t / c * 1e3} for k, (c, t) in sorted(prof.items())}}
    if hasattr(wl, "step_algorithmic_bytes"):
        a = wl.step_algorithmic_bytes()
        line["roofline_step"] = {"bound": "hbm", "achieved": a / (ms_local * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                 "frac": a / (ms_local * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes_per_step": a}
    elif hasattr(wl, "step_roofline"):
        line["roofline_step"] = wl.step_roofline(ms_local, peaks)
    line.update(extra)
    if world == 1 and cpu_baseline:
        line["cpu_baseline"] = wl.cpu_baseline(budget_s=cpu_budget_s)
        if hasattr(wl, "reference_ext"):
            line["reference_ext"] = wl.reference_ext()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default=os.environ.get("NGP_BENCH_WORKLOAD", DEFAULT_WORKLOAD), choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=0, help="override the per-GPU batch (development only)")
    ap.add_argument("--rays", type=int, default=0, help="nerf_train: rays per GPU per step (default: configs[1]/[2])")
    ap.add_argument("--pretrain", type=int, default=1024, help="nerf_train: untimed optimisation steps before the warm-up")
    ap.add_argument("--no-graph", action="store_true", help="nerf_train: launch the step eagerly instead of replaying CUDA graphs")
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--flush", action="store_true", help="flush L2 between steps even for the workloads whose per-step data exceeds L2")
    ap.add_argument("--adam-blocks", type=int, default=5, help="nerf_train: k_adam blocks per SM while it runs beside the marcher")
    ap.add_argument("--no-overlap", action="store_true", help="nerf_train: optimizer(n) strictly before march(n+1) (default: side by side on two streams)")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=INT", help="ngp_set_option before setup (kernel A/B tests; development only)")
    ap.add_argument("--sdf-chunks", type=int, default=0, help="sdf: level ranges of the pipelined table-gradient all-reduce (0 = trainer default: 4 when N > 1)")
    ap.add_argument("--sdf-bin-bits", type=int, default=-1, help="sdf: Morton binning of the batch, 0 = off, 5..7 (-1 = trainer default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--only", action="store_true", help="the named workload alone: no `workloads` block with the other BASELINE configs (N=1) "
                                                        "and no single-GPU leg at the global batch (N>1)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    wl = WORKLOADS[args.workload](args)
    if args.impl == "reference":
        rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
        run_reference(wl, args, rank, world)
        return

    import torch
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (the product path has no CPU fallback)"
    rank, world, local = dist_setup(args.gpus)
    dev = torch.device("cuda", local)
    import ngp_abi as abi
    peaks = load_peaks()
    for kv in args.opt:
        k, v = kv.split("=")
        abi.set_option(k, int(v))
    line = measure(wl, args, rank, world, dev, peaks, args.steps, args.warmup, cpu_baseline=not args.no_cpu_baseline)

    # N > 1 training: the scaling denominator SURVEY.md 8(d) config 3 names -- ONE GPU at the same global batch, same model state
    if world > 1 and args.workload == "nerf_train" and not args.only and not args.rays:
        single = wl.single_gpu_same_global_batch(rank, world, args.steps, args.warmup)
        if rank == 0:
            line["single_gpu_same_global_batch"] = single
            line["scaling_reference"] = "single_gpu_same_global_batch (strong scaling: the global batch is fixed at 2^18 rays/step)"

    # N = 1: the other BASELINE.json configs in the same line (configs[0] first: it leads BASELINE.json's metric)
    if world == 1 and not args.only and args.workload == DEFAULT_WORKLOAD:
        del wl
        torch.cuda.empty_cache()
        line["workloads"] = {}
        for name in ("hashgrid", "render", "sdf"):
            sub = WORKLOADS[name](args)
            sub_line = measure(sub, args, rank, world, dev, peaks, max(3, min(args.steps, 20)), 3, cpu_baseline=not args.no_cpu_baseline, cpu_budget_s=5.0)
            for k in ("n_gpus", "higher_is_better", "vs_baseline", "data", "clocks"):
                sub_line.pop(k, None)
            line["workloads"][name] = sub_line
            del sub
            torch.cuda.empty_cache()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
