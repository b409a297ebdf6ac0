"""Development helper: time the SDF step's table-gradient backward per level range (ngp_call_opts.level_begin / level_end),
to choose the ranges of the pipelined all-reduce (nerf/sdf_trainer.py).  python benchmarks/sdf_chunks.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "torch-ngp_b200"))
import torch
from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch

B = 1 << 20
tr = SDFTrainer(batch=B, use_graph=False)
p, y = sample_sdf_batch(B, "cuda", seed=1)
for _ in range(20):
    tr.train_step(p, y)
tr._forward_backward(grid_backward=False)
torch.cuda.synchronize()


def timed(ranges, reps=10):
    out = []
    for lb, le in ranges:
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); tr._grid_backward(lb, le); b.record(); b.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        ts.sort()
        out.append(round(ts[len(ts) // 2], 1))
    return out


for name, ranges in [("whole", [(0, 0)]), ("4 ranges", [(12, 16), (8, 12), (4, 8), (0, 4)]), ("2 ranges", [(8, 16), (0, 8)]),
                     ("12+4", [(4, 16), (0, 4)]), ("8+4+4", [(8, 16), (4, 8), (0, 4)]), ("4+12", [(12, 16), (0, 12)])]:
    t = timed(ranges)
    print(f"{name:10s} {ranges} us {t} sum {sum(t):.1f}", flush=True)
tr.g16.zero_()
