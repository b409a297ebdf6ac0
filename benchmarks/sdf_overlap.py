"""Development helper: timeline of the SDF step's deferred optimizer (side stream) against the binning of the next batch."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "torch-ngp_b200"))
import torch
from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch

B = 1 << 20
batches = [sample_sdf_batch(B, "cuda", seed=s) for s in range(2)]
for graph in (True, False):
    for cap in (0, 2, 4, 8):
        for ov in (True, False):
            tr = SDFTrainer(batch=B, use_graph=graph, overlap_optimizer=ov)
            tr.adam_blocks_per_sm = cap
            for i in range(10):
                tr.train_step(*batches[i % 2])
            tr.overlap_events = [] if ov else None
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for i in range(20):
                tr.train_step(*batches[i % 2])
            tr.flush()
            b.record(); b.synchronize()
            msg = f"graph={graph} adam_cap={cap} overlap={ov}: {a.elapsed_time(b) / 20 * 1e3:.1f} us/step"
            if ov:
                ev = tr.overlap_events[5:]
                f = lambda i, j: sum(e[i].elapsed_time(e[j]) for e in ev) / len(ev) * 1e3
                msg += f" | bin {f(0, 1):.1f} side {f(2, 3):.1f} side starts {f(0, 2):+.1f} after main, join {f(0, 4):.1f}"
            print(msg, flush=True)
            del tr
