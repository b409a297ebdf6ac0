import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import numpy as np, torch, ngp_abi as abi
from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
tr = SDFTrainer(batch=1 << 20, use_graph=False)
p, y = sample_sdf_batch(1 << 20, "cuda", seed=1)
for _ in range(5): tr.train_step(p, y)
def t(fn, iters=10):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3
def bwd():
    abi.check(abi.lib.ngp_grid_encode_backward(tr.g_enc.data_ptr(), tr.x01.data_ptr(), tr.table16.data_ptr(), tr.offsets.data_ptr(), tr.offsets_host.ctypes.data,
              tr.g16.data_ptr(), tr.B, 3, tr.C, tr.L, tr.S, tr.base_resolution, None, None, 0, 0, 0, abi.DTYPE_F16, abi.LAYOUT_BLC, abi.stream()))
for var in (2, 1, 0):
    for agg in (1, 0):
        abi.set_option("hashgrid_bwd_variant", var); abi.set_option("hashgrid_bwd_warp_agg", agg)
        print("bwd variant", var, "agg", agg, f"{t(bwd):.1f} us")
# sorted points (Morton order of the finest useful cell) -> runs of equal cells
x = tr.x01
key = ((x[:, 0] * 1023).long() << 20) | ((x[:, 1] * 1023).long() << 10) | (x[:, 2] * 1023).long()
order = torch.argsort(key)
tr.x01.copy_(x[order]); tr.g_enc.copy_(tr.g_enc[order])
abi.set_option("hashgrid_bwd_variant", 2); abi.set_option("hashgrid_bwd_warp_agg", 1)
print("bwd variant 2 agg 1, points sorted by cell:", f"{t(bwd):.1f} us")
