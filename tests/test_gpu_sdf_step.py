"""SDF training step (nerf/sdf_trainer.py, configs[4]) against the autograd path through the plugin modules
(GridEncoder + FFMLP + mape loss, i.e. sdf/netowrk_ff.py + sdf/utils.py::train_step) and against the CPU oracle step."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a.double() - b.double()).norm() / max(1e-30, b.double().norm()))


def _make(B=1 << 14, graph=False):
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    tr = SDFTrainer(batch=B, use_graph=graph, log2_hashmap_size=15, loss_scale=1024.0)
    p, y = sample_sdf_batch(B, "cuda", seed=1)
    return tr, p, y


def test_sdf_step_matches_autograd_path_and_oracle(oracle):
    import gridencoder, ffmlp
    from oracle import nerf_step_ref as R
    tr, p, y = _make()
    tr.points.copy_(p); tr.sdfs.copy_(y)
    enc = gridencoder.GridEncoder(num_levels=16, log2_hashmap_size=15, desired_resolution=2048).cuda()
    net = ffmlp.FFMLP(32, 1, 64, 3).cuda().train()
    with torch.no_grad():
        enc.embeddings.copy_(tr.embeddings); net.weights.copy_(tr.mlp_w)
    st = R.SDFRefState(tr.embeddings.cpu().numpy(), tr.mlp_w.cpu().numpy(), tr.offsets_host, tr.per_level_scale, tr.base_resolution, loss_scale=1024.0)
    tr._forward_backward()
    torch.cuda.synchronize()
    with torch.autocast("cuda", dtype=torch.half):
        pred = net(enc(p, bound=1))
    diff = (pred.float().squeeze(-1) - y).abs() / (y.abs() + 1e-2)
    loss = diff.mean()
    (loss * 1024.0).backward()
    assert abs(float(tr.state[7]) - float(loss.detach())) < 1e-3 * max(1.0, float(loss.detach()))
    assert _rel(tr.g32, net.weights.grad) < 2e-2
    assert _rel(tr.g16.float(), enc.embeddings.grad) < 2e-2
    loss_ref, grads = R.sdf_forward_backward(st, p.cpu().numpy(), y.cpu().numpy())
    assert abs(float(tr.state[7]) - loss_ref) < 1e-3 * max(1.0, loss_ref)
    assert _rel(tr.g32, torch.from_numpy(grads[1]).cuda()) < 2e-2
    assert _rel(tr.g16.float(), torch.from_numpy(grads[0]).cuda()) < 2e-2
    # optimizer incl. the MLP weight decay, fed with the device gradients
    R.sdf_adam_step(st, (tr.g16.float().cpu().numpy(), tr.g32.cpu().numpy().copy()))
    tr._optimizer_step()
    torch.cuda.synchronize()
    assert _rel(tr.mlp_w, torch.from_numpy(st.mlp_w).cuda()) < 1e-5
    assert _rel(tr.embeddings, torch.from_numpy(st.table).cuda()) < 1e-4


@pytest.mark.parametrize("graph", [False, True])
def test_sdf_training_converges(graph):
    from nerf.sdf_trainer import SDFTrainer, sample_sdf_batch
    tr = SDFTrainer(batch=1 << 16, use_graph=graph, log2_hashmap_size=17, lr=1e-3)
    batches = [sample_sdf_batch(1 << 16, "cuda", seed=s) for s in range(4)]
    losses = []
    for i in range(200):
        tr.train_step(*batches[i % 4])
        if i % 20 == 19:
            losses.append(tr.loss())
    assert np.isfinite(losses).all() and losses[-1] < 0.5 * losses[0], losses
    p, y = sample_sdf_batch(1 << 14, "cuda", seed=99)
    err = (tr.predict(p) - y).abs().mean()
    assert float(err) < 0.05, float(err)
