"""Development helper (torchrun, N >= 1): wall-clock ms/step of the SDF step with host or device batches, with or without the
per-step loss read-back, deferred optimizer on/off -- to see which of them interacts with the side-stream collective."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "torch-ngp_b200"))
import torch
import torch.distributed as dist
from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl")
B = 1 << 20
dev_b = [sample_sdf_batch(B, "cuda", seed=100 * rank + s) for s in range(4)]
host_b = [(p.cpu().pin_memory(), y.cpu().pin_memory()) for p, y in dev_b]


def sync():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


for ov in (False, True):
    tr = SDFTrainer(batch=B, world_size=world, overlap_optimizer=ov, grad_chunks=1)
    for i in range(8):
        tr.train_step(*dev_b[i % 4])
    for host in (False, True):
        for readback in (False, True):
            bs = host_b if host else dev_b
            get = None
            for i in range(4):
                tr.train_step(*bs[i % 4])
            sync()
            t0 = time.perf_counter()
            for i in range(30):
                tr.train_step(*bs[i % 4])
                if readback:
                    prev, get = get, tr.loss_async()
                    if prev is not None:
                        prev()
            if get is not None:
                get()
            sync()
            ms = (time.perf_counter() - t0) / 30 * 1e3
            if rank == 0:
                print(f"overlap={ov} host_batches={host} loss_readback={readback}: {ms:.3f} ms/step", flush=True)
    tr.flush()
    del tr
if world > 1:
    dist.destroy_process_group()
