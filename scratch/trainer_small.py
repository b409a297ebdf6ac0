import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import torch, nerf, ngp_abi as abi
from nerf import synthetic
mode = sys.argv[1] if len(sys.argv) > 1 else "default"
if mode == "split1": abi.set_option("ffmlp_bwd_split", 1)
if mode == "old": abi.set_option("ffmlp_bwd_variant", 0)
tr = nerf.NGPTrainer(bound=1.0, num_rays=4096, use_graph=False)
o, d, gt = synthetic.make_pool(4, 4096, "cuda", seed=3)
tr.set_pool(o, d, gt)
for i in range(3):
    tr.next_pool_batch(); tr.train_step()
    torch.cuda.synchronize()
    print("step", i, tr.loss(), int(tr.step_counter[0, 0]), tr.M, flush=True)
