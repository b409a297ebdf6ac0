"""Multi-GPU plumbing of the hot path: one process per GPU, torch.distributed (NCCL over NVLink; gloo in the CPU tests).

Training (BASELINE.json configs[2]): data-parallel over rays.  Every rank marches / shades its own N rays and holds a
full replica of the parameters; the only exchange is the gradient all-reduce -- one call for the fp16 hash-table
gradient (T*C halves) and one for the fp32 MLP gradients -- after which every rank runs the same fused Adam with
grad_div = world_size (sum of per-rank means / world = mean over the global batch).
Inference (configs[3]): image rows are split over the ranks, no collective: contiguous bands (row_band) or interleaved
stripes (row_stripes) -- stripes balance the load, since the rows that cross the object cost ~10x the rows of empty sky
(measured on 8 B200: bands 3.6x one GPU, stripes see BASELINE.md).
"""
import torch


def allreduce_gradients(g16, g32, group=None):
    """Sum the flat gradient buffers over the ranks, in place.  g16: fp16 table gradient, g32: fp32 MLP gradients."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    dist.all_reduce(g16.view(-1), group=group)
    dist.all_reduce(g32.view(-1), group=group)


def row_band(height, rank, world):
    """Rows [r0, r1) of an image that rank renders: contiguous bands, sizes differ by at most one row."""
    base, extra = divmod(height, world)
    r0 = rank * base + min(rank, extra)
    return r0, r0 + base + (1 if rank < extra else 0)


def row_stripes(height, rank, world, stripe=8):
    """Row indices (ascending int64 tensor) that rank renders: stripes of `stripe` rows dealt round-robin to the ranks.
    Disjoint over the ranks and covering [0, height)."""
    rows = torch.arange(height, dtype=torch.int64)
    return rows[(rows // stripe) % world == rank]


def rays_per_rank(global_rays, world):
    assert global_rays % world == 0, "the global ray batch must divide evenly over the ranks"
    return global_rays // world
