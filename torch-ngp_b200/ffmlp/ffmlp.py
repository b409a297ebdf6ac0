"""ffmlp.FFMLP / ffmlp_forward -- same public API as /root/reference/ffmlp/ffmlp.py (`_ffmlp_forward` :15-86,
`FFMLP` :99-168), backed by the tcgen05 kernels of libngp_b200.so.

Same constraints, parameter layout and seeded initialisation as the reference (one flat fp32 `weights`
Parameter: layer 0 [hidden, in], num_layers-1 x [hidden, hidden], last [16, hidden], each row-major
[out, in]; manual_seed(42), U(+-sqrt(3/hidden))).  Differences, all below the API:
  * no batch padding copy (ffmlp.py:157-159 cats up to 128 zero rows): the kernels mask the ragged tail;
  * no forward_buffer / backward_buffer for hidden_dim <= 128: activations stay in shared memory / TMEM and the
    backward recomputes the forward (480 -> 96 B/point forward, ~2100 -> 160 B/point backward); hidden_dim 256
    keeps both buffers in the reference's layout (csrc/ffmlp_wide.cu);
  * fp32 accumulation (TMEM) and fp32 weight gradients; weights are read as fp32 and rounded to fp16 while
    staged into shared memory, so there is no per-call cast of the parameter vector.
"""
import math

import torch
import torch.nn as nn
from torch.autograd import Function

import ngp_abi as abi
from . import backend as _backend


def available():
    """True when libngp_b200.so carries the tcgen05 FFMLP kernels."""
    return abi.get_option("ffmlp_available") != 0


class _ffmlp_forward(Function):
    @staticmethod
    def forward(ctx, inputs, weights, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                inference=False, calc_grad_inputs=False):
        # the reference casts every floating input to half (custom_fwd(cast_inputs=torch.half), ffmlp.py:18);
        # here only the activations are cast -- the kernel rounds the fp32 weights itself.
        B = inputs.shape[0]
        inputs = inputs.to(torch.half).contiguous()
        weights = weights.contiguous()
        outputs = torch.empty(B, output_dim, device=inputs.device, dtype=inputs.dtype)
        with torch.amp.autocast("cuda", enabled=False):
            if not inference:
                # hidden_dim 256 keeps the reference's forward_buffer (ffmlp.py:33); narrower nets recompute in the backward
                fb = torch.empty(num_layers, B, hidden_dim, device=inputs.device, dtype=inputs.dtype) if hidden_dim == 256 else None
                _backend.ffmlp_forward(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                       output_activation, fb, outputs)
                ctx.save_for_backward(inputs, weights, *([fb] if fb is not None else []))
                ctx.dims = (input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs)
            else:
                _backend.ffmlp_inference(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                         output_activation, None, outputs)
        return outputs

    @staticmethod
    def backward(ctx, grad):
        B = grad.shape[0]
        grad = grad.to(torch.half).contiguous()
        inputs, weights, *rest = ctx.saved_tensors
        input_dim, output_dim, hidden_dim, num_layers, activation, output_activation, calc_grad_inputs = ctx.dims
        grad_inputs = torch.empty_like(inputs) if calc_grad_inputs else None
        grad_weights = torch.empty_like(weights)  # fp32 for the fp32 Parameter; overwritten by the kernel
        fb = rest[0] if rest else None
        bb = torch.empty_like(fb) if fb is not None else None  # ffmlp.py:70 (every row in use is written before it is read)
        _backend.ffmlp_backward(grad, inputs, weights, fb, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                                output_activation, calc_grad_inputs, bb, grad_inputs, grad_weights)
        return grad_inputs, grad_weights, None, None, None, None, None, None, None, None


ffmlp_forward = _ffmlp_forward.apply


def convert_activation(act):
    return {"relu": 0, "exponential": 1, "sine": 2, "sigmoid": 3, "squareplus": 4, "softplus": 5}.get(act, 6)


class FFMLP(nn.Module):
    def __init__(self, input_dim, output_dim, hidden_dim, num_layers, activation="relu"):
        super().__init__()
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.hidden_dim = hidden_dim
        self.num_layers = num_layers
        self.activation = convert_activation(activation)
        self.output_activation = convert_activation("none")  # not supported currently (as the reference)
        self.tensorcore_width = 16

        assert hidden_dim in [16, 32, 64, 128, 256], f"FFMLP only support hidden_dim in [16, 32, 64, 128, 256], but got {hidden_dim}"
        assert input_dim > 0 and input_dim % 16 == 0, f"FFMLP input_dim should be 16 * m (m  > 0), but got {input_dim}"
        assert output_dim <= 16, f"FFMLP current only supports output dim <= 16, but got {output_dim}"
        assert num_layers >= 2, f"FFMLP num_layers should be larger than 2 (3 matmuls), but got {num_layers}"

        self.padded_output_dim = int(math.ceil(output_dim / 16)) * 16
        self.num_parameters = hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + self.padded_output_dim)
        self.weights = nn.Parameter(torch.zeros(self.num_parameters))
        self.reset_parameters()
        _backend.allocate_splitk(self.num_layers + 1)  # no-op: there is no split-k stream pool here

    def cleanup(self):
        _backend.free_splitk()

    def __repr__(self):
        return (f"FFMLP: input_dim={self.input_dim} output_dim={self.output_dim} hidden_dim={self.hidden_dim} "
                f"num_layers={self.num_layers} activation={self.activation}")

    def reset_parameters(self):
        torch.manual_seed(42)
        std = math.sqrt(3 / self.hidden_dim)
        self.weights.data.uniform_(-std, std)

    def forward(self, inputs):
        # inputs [B, input_dim] -> [B, output_dim] (half)
        B, C = inputs.shape
        outputs = ffmlp_forward(inputs, self.weights, self.input_dim, self.padded_output_dim, self.hidden_dim, self.num_layers,
                                self.activation, self.output_activation, not self.training, inputs.requires_grad)
        if self.padded_output_dim != self.output_dim:
            outputs = outputs[:, :self.output_dim]
        return outputs
