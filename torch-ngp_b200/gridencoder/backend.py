"""`_backend` shim: the reference's pybind11 module `_gridencoder` (gridencoder/src/bindings.cpp:5-9),
same three names and the same positional arguments, forwarding to the C ABI of libngp_b200.so."""
import numpy as np
import torch

import ngp_abi as abi

_offsets_cache = {}


def host_offsets(offsets):
    """Host copy of the (tiny, immutable) level-offset table; one D2H read per distinct tensor."""
    key = (offsets.data_ptr(), offsets._version, offsets.numel())
    hit = _offsets_cache.get(key)
    if hit is None:
        if len(_offsets_cache) > 64:
            _offsets_cache.clear()
        hit = np.ascontiguousarray(offsets.detach().cpu().numpy().astype(np.int32))
        _offsets_cache[key] = hit
    return hit


def _check_common(inputs, embeddings, offsets):
    abi.check_tensor(inputs, "inputs")
    abi.check_tensor(embeddings, "embeddings")
    abi.check_tensor(offsets, "offsets", torch.int32, "an int")
    if inputs.dtype != torch.float32:
        # the reference only checks "floating" and then reinterprets as float (gridencoder.cu:458,466)
        raise RuntimeError("inputs must be a float32 tensor")


def grid_encode_forward(inputs, embeddings, offsets, outputs, B, D, C, L, S, H, dy_dx, gridtype, align_corners, interp=0,
                        layout=abi.LAYOUT_LBC, offsets_host=None, opts=None):
    """Reference positional arguments (gridencoder.h:12).  Keyword extras of this library: layout (NGP_LAYOUT_*),
    offsets_host (skips the cached D2H read), opts (abi.CallOpts: per-call options)."""
    _check_common(inputs, embeddings, offsets)
    abi.check_tensor(outputs, "outputs", embeddings.dtype, "the embeddings' dtype", B * L * C)
    abi.check_numel(inputs, B * D, "inputs")
    abi.check_numel(offsets, L + 1, "offsets")
    if dy_dx is not None:
        abi.check_tensor(dy_dx, "dy_dx", embeddings.dtype, "the embeddings' dtype", B * L * D * C)
    oh = host_offsets(offsets) if offsets_host is None else offsets_host
    abi.check(abi.lib.ngp_grid_encode_forward(
        inputs.data_ptr(), embeddings.data_ptr(), offsets.data_ptr(), oh.ctypes.data, outputs.data_ptr(), B, D, C, L,
        float(S), H, abi.ptr(dy_dx), gridtype, int(bool(align_corners)), interp, abi.dtype_code(embeddings, "embeddings"),
        layout, abi.stream(), opts))


def grid_encode_backward(grad, inputs, embeddings, offsets, grad_embeddings, B, D, C, L, S, H, dy_dx, grad_inputs, gridtype,
                         align_corners, interp=0, layout=abi.LAYOUT_LBC, offsets_host=None, opts=None):
    _check_common(inputs, embeddings, offsets)
    abi.check_tensor(grad, "grad", None, None, B * L * C)
    abi.check_tensor(grad_embeddings, "grad_embeddings", grad.dtype, "the grad's dtype", embeddings.numel())
    if (dy_dx is None) != (grad_inputs is None):
        raise RuntimeError("dy_dx and grad_inputs must be given together")
    if dy_dx is not None:
        abi.check_tensor(dy_dx, "dy_dx", grad.dtype, "the grad's dtype", B * L * D * C)
        abi.check_tensor(grad_inputs, "grad_inputs", grad.dtype, "the grad's dtype", B * D)
    oh = host_offsets(offsets) if offsets_host is None else offsets_host
    abi.check(abi.lib.ngp_grid_encode_backward(
        grad.data_ptr(), inputs.data_ptr(), embeddings.data_ptr(), offsets.data_ptr(), oh.ctypes.data,
        grad_embeddings.data_ptr(), B, D, C, L, float(S), H, abi.ptr(dy_dx), abi.ptr(grad_inputs), gridtype,
        int(bool(align_corners)), interp, abi.dtype_code(grad, "grad"), layout, abi.stream(), opts))


def grad_total_variation(inputs, embeddings, grad, offsets, weight, B, D, C, L, S, H, gridtype, align_corners):
    abi.check_tensor(inputs, "inputs", embeddings.dtype, "the embeddings' dtype", B * D)
    abi.check_tensor(embeddings, "embeddings")
    abi.check_tensor(grad, "grad", embeddings.dtype, "the embeddings' dtype", embeddings.numel())
    abi.check_tensor(offsets, "offsets", torch.int32, "an int", L + 1)
    abi.check(abi.lib.ngp_grad_total_variation(
        inputs.data_ptr(), embeddings.data_ptr(), grad.data_ptr(), offsets.data_ptr(), float(weight), B, D, C, L, float(S),
        H, gridtype, int(bool(align_corners)), abi.dtype_code(embeddings, "embeddings"), abi.stream()))
