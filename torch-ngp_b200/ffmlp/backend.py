"""`_backend` shim: the reference's pybind11 module `_ffmlp` (ffmlp/src/bindings.cpp:5-10), same five
names and the same positional arguments, forwarding to the C ABI of libngp_b200.so.

Differences below the API (include/ngp_b200.h): `weights` may be fp32 (converted to fp16 while staged
into shared memory, so no per-call cast kernel) or fp16; `forward_buffer` / `backward_buffer` /
`inference_buffer` may be None for hidden_dim <= 128 -- the tcgen05 kernels keep every activation in shared memory /
TMEM and recompute the forward inside the backward instead of round-tripping [num_layers, B, hidden] through HBM
(hidden_dim 256 uses both buffers in the reference's layout: neither its weights nor its weight-gradient accumulators
fit on the SM next to a batch tile);
`grad_weights` may be fp32 (no K = B reduction in half precision)."""
import torch

import ngp_abi as abi

_f16 = torch.float16
_ws_cache = {}


def _workspace(device, nbytes):
    """Zero-initialised dW accumulation workspace, one per (device, stream): the kernel leaves it zeroed for the next
    call ON THE SAME STREAM; two streams (or two threads with their own streams) never share one."""
    key = (device.type, device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 16), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def _check_dims(B, input_dim, output_dim, hidden_dim, num_layers):
    if hidden_dim not in (16, 32, 64, 128, 256):
        raise RuntimeError("hidden_dim should in [16, 32, 64, 128, 256]")


def _n_weights(input_dim, output_dim, hidden_dim, num_layers):
    return hidden_dim * (input_dim + hidden_dim * (num_layers - 1) + output_dim)


def _wdtype(weights):
    if weights.dtype not in (torch.float16, torch.float32):
        raise RuntimeError("weights must be a half or float tensor")
    return abi.dtype_code(weights, "weights")


def ffmlp_forward(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                  forward_buffer, outputs, opts=None):
    _check_dims(B, input_dim, output_dim, hidden_dim, num_layers)
    abi.check_tensor(inputs, "inputs", _f16, "a half", B * input_dim)
    abi.check_tensor(weights, "weights", None, None, _n_weights(input_dim, output_dim, hidden_dim, num_layers))
    abi.check_tensor(outputs, "outputs", _f16, "a half", B * output_dim)
    if forward_buffer is not None and forward_buffer.numel() > 0:
        abi.check_tensor(forward_buffer, "forward_buffer", _f16, "a half", num_layers * B * hidden_dim)
    else:
        forward_buffer = None
    abi.check(abi.lib.ngp_ffmlp_forward(inputs.data_ptr(), weights.data_ptr(), _wdtype(weights), B, input_dim, output_dim, hidden_dim,
                                        num_layers, activation, output_activation, abi.ptr(forward_buffer), outputs.data_ptr(),
                                        abi.stream(), opts))


def ffmlp_inference(inputs, weights, B, input_dim, output_dim, hidden_dim, num_layers, activation, output_activation,
                    inference_buffer, outputs, opts=None):
    _check_dims(B, input_dim, output_dim, hidden_dim, num_layers)
    abi.check_tensor(inputs, "inputs", _f16, "a half", B * input_dim)
    abi.check_tensor(weights, "weights", None, None, _n_weights(input_dim, output_dim, hidden_dim, num_layers))
    abi.check_tensor(outputs, "outputs", _f16, "a half", B * output_dim)
    abi.check(abi.lib.ngp_ffmlp_inference(inputs.data_ptr(), weights.data_ptr(), _wdtype(weights), B, input_dim, output_dim, hidden_dim,
                                          num_layers, activation, output_activation, None, outputs.data_ptr(), abi.stream(), opts))


def ffmlp_backward(grad, inputs, weights, forward_buffer, B, input_dim, output_dim, hidden_dim, num_layers, activation,
                   output_activation, calc_grad_inputs, backward_buffer, grad_inputs, grad_weights, opts=None):
    _check_dims(B, input_dim, output_dim, hidden_dim, num_layers)
    nw = _n_weights(input_dim, output_dim, hidden_dim, num_layers)
    abi.check_tensor(grad, "grad", _f16, "a half", B * output_dim)
    abi.check_tensor(inputs, "inputs", _f16, "a half", B * input_dim)
    abi.check_tensor(weights, "weights", None, None, nw)
    abi.check_tensor(grad_weights, "grad_weights", None, None, nw)
    if grad_weights.dtype not in (torch.float16, torch.float32):
        raise RuntimeError("grad_weights must be a half or float tensor")
    if calc_grad_inputs:
        abi.check_tensor(grad_inputs, "grad_inputs", _f16, "a half", B * input_dim)
    nbytes = abi.lib.ngp_ffmlp_backward_workspace_bytes(input_dim, output_dim, hidden_dim, num_layers)
    ws = _workspace(grad.device, nbytes)
    fb = bb = None
    if hidden_dim == 256:  # the one width that reads them (two-phase backward, csrc/ffmlp_wide.cu); narrower nets recompute on the SM
        if forward_buffer is None or backward_buffer is None:
            raise RuntimeError("ffmlp_backward: hidden_dim 256 needs forward_buffer and backward_buffer ([num_layers, B, 256] half)")
        abi.check_tensor(forward_buffer, "forward_buffer", _f16, "a half", num_layers * B * hidden_dim)
        abi.check_tensor(backward_buffer, "backward_buffer", _f16, "a half", num_layers * B * hidden_dim)
        fb, bb = forward_buffer.data_ptr(), backward_buffer.data_ptr()
    abi.check(abi.lib.ngp_ffmlp_backward(
        grad.data_ptr(), inputs.data_ptr(), weights.data_ptr(), _wdtype(weights), fb, B, input_dim, output_dim, hidden_dim, num_layers,
        activation, output_activation, int(bool(calc_grad_inputs)), bb, grad_inputs.data_ptr() if calc_grad_inputs else None,
        grad_weights.data_ptr(), abi.dtype_code(grad_weights, "grad_weights"), ws.data_ptr(), ws.numel(), abi.stream(), opts))


def allocate_splitk(size):
    abi.check(abi.lib.ngp_allocate_splitk(size))


def free_splitk():
    abi.check(abi.lib.ngp_free_splitk())
