"""Host -> device input feed of the trainers: the batch of step n+1 crosses PCIe on a side stream while the GPU still computes
step n.  The reference's loop does `data = next(loader); rays = data['rays_o'] ...` with tensors its provider already put on
the GPU (nerf/provider.py:182-206 preloads the whole dataset); a caller that keeps its data in pinned host memory gets the same
effect here without a blocking copy in front of every step."""
import torch


class HostFeed:
    """stage(dsts, srcs): every src that lives on the host is copied into a staging buffer on a side stream (the host has run
    ahead of the GPU, so that copy overlaps the previous step's kernels); the current stream then waits for the copy and moves
    staging -> dst with device-to-device copies.  Two staging sets alternate, guarded by events, so a set is never overwritten
    while the step that reads it is still queued.  Device-resident srcs are copied directly.  The host tensors must stay
    unchanged until their copy has run (as with any non_blocking copy from pinned memory)."""

    def __init__(self, device, depth=2):
        self.dev = torch.device(device)
        self.stream = torch.cuda.Stream(self.dev)
        self.depth, self.i = depth, 0
        self.staging = [None] * depth
        self.copied = [torch.cuda.Event() for _ in range(depth)]
        self.consumed = [None] * depth

    def stage(self, dsts, srcs, after=None):
        """after: an event the host -> device copy also waits for (SDFTrainer: the end of the previous step's gradient all-reduce --
        a copy in flight while the collective runs stretched the step by 0.16 ms on 2 GPUs, measured with benchmarks/sdf_e2e_probe.py)."""
        if all(s.is_cuda for s in srcs):
            for d, s in zip(dsts, srcs):
                d.copy_(s, non_blocking=True)
            return
        i = self.i = (self.i + 1) % self.depth
        if self.staging[i] is None or any(a.shape != d.shape or a.dtype != d.dtype for a, d in zip(self.staging[i], dsts)):
            self.staging[i] = [torch.empty_like(d) for d in dsts]
        main = torch.cuda.current_stream(self.dev)
        if self.consumed[i] is not None:
            self.stream.wait_event(self.consumed[i])  # the step that read this staging set two calls ago
        if after is not None:
            self.stream.wait_event(after)
        with torch.cuda.stream(self.stream):
            for a, s in zip(self.staging[i], srcs):
                a.copy_(s.reshape(a.shape) if s.shape != a.shape else s, non_blocking=True)
            self.copied[i].record(self.stream)
        main.wait_event(self.copied[i])
        for d, a in zip(dsts, self.staging[i]):
            d.copy_(a, non_blocking=True)
        ev = self.consumed[i] or torch.cuda.Event()
        ev.record(main)
        self.consumed[i] = ev
