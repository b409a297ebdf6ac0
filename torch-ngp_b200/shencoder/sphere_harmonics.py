"""shencoder.SHEncoder -- same public API as /root/reference/shencoder/sphere_harmonics.py (:13-87), degree 1..4,
forward only (view directions are data, not parameters; the reference computes dy_dx only when the input requires
grad, which its NeRF networks never ask for)."""
import torch
import torch.nn as nn
from torch.autograd import Function
from torch.amp import custom_fwd

import ngp_abi as abi


class _sh_encoder(Function):
    @staticmethod
    @custom_fwd(device_type="cuda", cast_inputs=torch.float32)  # as the reference: fp32 in, fp32 out, even under autocast
    def forward(ctx, inputs, degree, calc_grad_inputs=False):
        if calc_grad_inputs:
            raise NotImplementedError("SHEncoder: gradients w.r.t. the directions are outside the accelerated path")
        if not inputs.is_cuda:
            raise RuntimeError("SHEncoder needs CUDA tensors (there is no CPU path)")
        inputs = inputs.contiguous()
        B = inputs.shape[0]
        outputs = torch.empty(B, degree ** 2, dtype=torch.float32, device=inputs.device)
        abi.check(abi.lib.ngp_sh_encode_forward(inputs.data_ptr(), outputs.data_ptr(), B, degree, degree ** 2, abi.DTYPE_F32, abi.stream()))
        return outputs


sh_encode = _sh_encoder.apply


class SHEncoder(nn.Module):
    def __init__(self, input_dim=3, degree=4):
        super().__init__()
        assert input_dim == 3, "SH encoder only support input dim == 3"
        assert 1 <= degree <= 4, "this build carries SH degrees 1..4"
        self.input_dim, self.degree, self.output_dim = input_dim, degree, degree ** 2

    def __repr__(self):
        return f"SHEncoder: input_dim={self.input_dim} degree={self.degree}"

    def forward(self, inputs, size=1):
        inputs = inputs / size
        prefix = inputs.shape[:-1]
        out = sh_encode(inputs.reshape(-1, self.input_dim), self.degree, inputs.requires_grad)
        return out.reshape(*prefix, self.output_dim)
