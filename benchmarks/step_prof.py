"""Two eager training steps of configs[1] between cudaProfilerStart/Stop, after an untimed pre-training: the target of
`ncu --profile-from-start off` captures (profiles/).  Development tool; bench.py is the contract."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "torch-ngp_b200"))
import torch  # noqa: E402


def main():
    import nerf
    from nerf import synthetic
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    torch.manual_seed(1000)
    tr = nerf.NGPTrainer(bound=1.0, num_rays=4096, device="cuda", seed=0, use_graph=False, overlap_optimizer=False)
    o, d, gt = synthetic.make_pool(16, 4096, "cuda", seed=1000)
    tr.set_pool(o, d, gt)
    for _ in range(1025):  # the occupancy grid (and the samples per ray) of a scene in training, as bench.py pre-trains
        tr.next_pool_batch()
        tr.train_step()
    torch.cuda.synchronize()
    tr.update_interval = 0
    torch.cuda.profiler.start()
    for _ in range(steps):
        tr.next_pool_batch()
        tr.train_step()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("samples", int(tr.step_counter[0, 0]), "loss", tr.loss())


if __name__ == "__main__":
    main()
