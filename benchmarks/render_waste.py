"""How many of the sample rows the inference loop shades are real samples (delta > 0)?  Development probe."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import torch
import nerf
from nerf import synthetic
import ngp_abi as abi

tr = nerf.NGPTrainer(bound=1.0, num_rays=4096, device="cuda", seed=0, use_graph=True)
o, d, gt = synthetic.make_pool(16, 4096, "cuda", seed=1)
tr.set_pool(o, d, gt)
for _ in range(1024):
    tr.next_pool_batch(); tr.train_step()
W, H = 1920, 1080
ro, rd = synthetic.frame_rays(W, H, "cuda", pose_index=7)
real = torch.zeros(1, dtype=torch.int64, device="cuda")
rows = torch.zeros(1, dtype=torch.int64, device="cuda")
orig = tr._render_iteration
def probe(rs, n_alive, n_step, cur, noises, T_thresh):
    orig(rs, n_alive, n_step, cur, noises, T_thresh)
    M = n_alive * n_step
    n = int(rs["rows_dev"].item())
    real.add_((rs["deltas"][:n, 0] > 0).sum()); rows.add_(n)
tr._render_iteration = probe
for lohi in ((2, 64), (1, 8)):
    real.zero_(); rows.zero_()
    tr.render(ro, rd, use_graph=False, n_step_range=lohi, sync_every=1)
    torch.cuda.synchronize()
    print(lohi, "rows shaded", int(rows), "real samples", int(real), "fraction", float(real) / float(rows), "iters", tr.last_render_iters)
print("occupied cells", int((tr.density_grid > min(tr.mean_density(), tr.density_thresh)).sum()), "of", tr.density_grid.numel())
