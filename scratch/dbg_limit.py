import sys; sys.path.insert(0, 'torch-ngp_b200')
import torch, ngp_abi as abi, ffmlp
torch.manual_seed(0)
B = 5000
net = ffmlp.FFMLP(32, 16, 64, 2).cuda().train()
x = (torch.randn(B, 32, device='cuda') * 0.5).half()
g = torch.randn(B, 16, device='cuda').half()
cnt = torch.zeros(2, dtype=torch.int32, device='cuda')
def run(xs, gs):
    net.weights.grad = None
    xs = xs.clone().requires_grad_(True)
    y = net(xs); y.backward(gs)
    return net.weights.grad.clone(), xs.grad.clone()
for n in (5000, 4993, 4900, 4864, 4000, 3217, 128, 1):
    ref, dx_ref = run(x[:n].contiguous(), g[:n].contiguous())
    cnt[0] = n
    abi.lib.ngp_set_batch_limit(cnt.data_ptr())
    got, dx = run(x, g)
    abi.lib.ngp_set_batch_limit(None)
    torch.cuda.synchronize()
    print(n, float((got - ref).abs().max()), float(ref.abs().max()), bool(torch.isnan(got).any()), float((dx[:n].float() - dx_ref.float()).abs().max()))
