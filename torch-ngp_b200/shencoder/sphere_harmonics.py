"""shencoder.SHEncoder -- same public API as /root/reference/shencoder/sphere_harmonics.py (:13-87), degree 1..8.
Gradients w.r.t. the directions (only when the input requires grad, as in the reference) come from one fused kernel that
contracts the upstream gradient with the analytic derivatives -- no [B, degree^2 * 3] dy_dx buffer."""
import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_bwd, custom_fwd

import ngp_abi as abi


class _sh_encoder(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)  # as the reference: fp32 in, fp32 out, even under autocast
    def forward(ctx, inputs, degree, calc_grad_inputs=False):
        if not inputs.is_cuda:
            raise RuntimeError("SHEncoder needs CUDA tensors (there is no CPU path)")
        inputs = inputs.contiguous()
        B = inputs.shape[0]
        outputs = torch.empty(B, degree ** 2, dtype=torch.float32, device=inputs.device)
        abi.check(abi.lib.ngp_sh_encode_forward(inputs.data_ptr(), outputs.data_ptr(), B, degree, degree ** 2, abi.DTYPE_F32, abi.stream()))
        ctx.degree, ctx.calc_grad_inputs = degree, bool(calc_grad_inputs)
        if calc_grad_inputs:
            ctx.save_for_backward(inputs)
        return outputs

    @staticmethod
    @custom_bwd(device_type="cuda")
    def backward(ctx, grad):
        if not ctx.calc_grad_inputs:
            return None, None, None
        (inputs,) = ctx.saved_tensors
        grad = grad.float().contiguous()
        grad_inputs = torch.empty_like(inputs)
        abi.check(abi.lib.ngp_sh_encode_backward(grad.data_ptr(), inputs.data_ptr(), grad_inputs.data_ptr(), inputs.shape[0], ctx.degree,
                                                 ctx.degree ** 2, abi.stream()))
        return grad_inputs, None, None


sh_encode = _sh_encoder.apply


class SHEncoder(nn.Module):
    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        assert input_dim == 3, "SH encoder only support input dim == 3"
        assert 1 <= degree <= 8, "SH encoder only supports degree in [1, 8]"
        self.input_dim, self.degree, self.output_dim = input_dim, degree, degree ** 2

    def __repr__(self):
        return f"SHEncoder: input_dim={self.input_dim} degree={self.degree}"

    def forward(self, inputs, size=1):
        inputs = inputs / size
        prefix = inputs.shape[:-1]
        out = sh_encode(inputs.reshape(-1, self.input_dim), self.degree, inputs.requires_grad)
        return out.reshape(*prefix, self.output_dim)
