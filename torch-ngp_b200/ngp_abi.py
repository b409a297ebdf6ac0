"""ctypes binding of libngp_b200.so (C ABI declared in include/ngp_b200.h).

This is the only place where Python touches the native library.  PyTorch is used for device
memory and streams only: every call passes raw ``tensor.data_ptr()`` values, sizes and the
current CUDA stream.  There is NO CPU fallback: if the library is missing (not built) the import
fails loudly, and every op refuses non-CUDA tensors.
"""
import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NGP_B200_LIB") or os.path.join(_HERE, "lib", "libngp_b200.so")  # override: developer builds (benchmarks/)

DTYPE_F32, DTYPE_F16, DTYPE_F64 = 0, 1, 2
LAYOUT_LBC, LAYOUT_BLC = 0, 1

_u32, _f32, _int, _sz, _vp = ctypes.c_uint32, ctypes.c_float, ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p

OPT_EXACT_INPUT_GRAD, OPT_GRID_BWD_SMEM_COARSE, OPT_GRID_BWD_GROUP_RED = 1, 2, 4
COMPACT_STATIC_MAX_N = 4194304  # NGP_COMPACT_STATIC_MAX_N


class CallOpts(ctypes.Structure):
    """ngp_call_opts (include/ngp_b200.h): per-call options, passed by pointer as the last argument of the per-sample
    entry points (None = defaults).  Nothing here is process-wide."""
    _fields_ = [("batch_limit", _vp), ("input_bound", _f32), ("flags", _u32), ("grid_cap_blocks_per_sm", ctypes.c_int32),
                ("level_begin", _u32), ("level_end", _u32)]


_op = ctypes.POINTER(CallOpts)


def call_opts(batch_limit=None, input_bound=0.0, flags=0, grid_cap_blocks_per_sm=0, level_begin=0, level_end=0):
    """-> a CallOpts to pass as `opts` (ctypes converts it to a pointer); batch_limit: int32 device tensor or None."""
    return CallOpts(None if batch_limit is None else batch_limit.data_ptr(), float(input_bound), int(flags), int(grid_cap_blocks_per_sm),
                    int(level_begin), int(level_end))

# name -> (restype, argtypes): mirrors include/ngp_b200.h one to one (tests/test_abi.py checks the
# header against this table and against the symbols the .so exports).
SIGNATURES = {
    "ngp_last_error": (ctypes.c_char_p, []),
    "ngp_abi_version": (_int, []),
    "ngp_set_option": (_int, [ctypes.c_char_p, _int]),
    "ngp_get_option": (_int, [ctypes.c_char_p, ctypes.POINTER(_int)]),
    "ngp_launch_count": (ctypes.c_longlong, []),
    "ngp_profile_collect": (_int, [ctypes.c_char_p, _sz]),
    "ngp_grid_encode_forward": (_int, [_vp, _vp, _vp, _vp, _vp, _u32, _u32, _u32, _u32, _f32, _u32, _vp, _u32, _int, _u32, _int, _int, _vp, _op]),
    "ngp_grid_encode_backward": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _u32, _u32, _u32, _u32, _f32, _u32, _vp, _vp, _u32, _int, _u32, _int, _int, _vp, _op]),
    "ngp_grad_total_variation": (_int, [_vp, _vp, _vp, _vp, _f32, _u32, _u32, _u32, _u32, _f32, _u32, _u32, _int, _int, _vp]),
    "ngp_near_far_from_aabb": (_int, [_vp, _vp, _vp, _u32, _f32, _vp, _vp, _vp]),
    "ngp_sph_from_ray": (_int, [_vp, _vp, _f32, _u32, _vp, _vp]),
    "ngp_morton3D": (_int, [_vp, _u32, _vp, _vp]),
    "ngp_morton3D_invert": (_int, [_vp, _u32, _vp, _vp]),
    "ngp_mark_untrained_grid": (_int, [_vp, _u32, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_double, _f32, _u32, _u32, _vp, _vp, _vp]),
    "ngp_packbits": (_int, [_vp, _u32, _f32, _vp, _vp]),
    "ngp_march_rays_train_workspace_bytes": (_sz, [_u32]),
    "ngp_march_rays_train_workspace_bytes_ex": (_sz, [_u32, _f32, _u32]),
    "ngp_march_rays_train": (_int, [_vp, _vp, _vp, _f32, _f32, _u32, _u32, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ngp_composite_rays_train_forward": (_int, [_vp, _vp, _vp, _vp, _u32, _u32, _f32, _vp, _vp, _vp, _vp]),
    "ngp_composite_rays_train_backward": (_int, [_vp] * 10 + [_u32, _u32, _f32, _vp, _vp, _vp]),
    "ngp_march_rays": (_int, [_u32, _vp, _u32, _vp, _vp, _vp, _vp, _f32, _f32, _u32, _u32, _u32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ngp_composite_rays": (_int, [_u32, _vp, _u32, _f32, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ngp_compact_alive_workspace_bytes": (_sz, [_u32]),
    "ngp_compact_alive": (_int, [_vp, _u32, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ngp_compact_alive_static": (_int, [_vp, _u32, _vp, _vp, _vp, _vp]),
    "ngp_ffmlp_forward": (_int, [_vp, _vp, _int, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _vp, _vp, _op]),
    "ngp_ffmlp_inference": (_int, [_vp, _vp, _int, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp, _vp, _vp, _op]),
    "ngp_ffmlp_backward": (_int, [_vp, _vp, _vp, _int, _vp, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _int, _vp, _vp, _vp, _int, _vp, _sz, _vp, _op]),
    "ngp_ffmlp_backward_workspace_bytes": (_sz, [_u32, _u32, _u32, _u32]),
    "ngp_allocate_splitk": (_int, [_sz]),
    "ngp_free_splitk": (_int, []),
    "ngp_sh_encode_forward": (_int, [_vp, _vp, _u32, _u32, _u32, _int, _vp]),
    "ngp_get_rays": (_int, [_vp, _u32, _f32, _f32, _f32, _f32, _u32, _vp, _u32, _u32, _vp, _vp, _vp]),
    "ngp_sh_encode_backward": (_int, [_vp, _vp, _vp, _u32, _u32, _u32, _vp]),
    "ngp_nerf_x01": (_int, [_vp, _f32, _vp, _u32, _vp, _op]),
    "ngp_nerf_mid_forward": (_int, [_vp, _vp, _f32, _vp, _vp, _u32, _vp, _op]),
    "ngp_nerf_rgb_forward": (_int, [_vp, _vp, _u32, _vp, _op]),
    "ngp_nerf_loss": (_int, [_vp, _vp, _vp, _f32, _u32, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ngp_nerf_rgb_backward": (_int, [_vp, _vp, _vp, _u32, _vp, _op]),
    "ngp_nerf_mid_backward": (_int, [_vp, _vp, _vp, _f32, _vp, _u32, _vp, _op]),
    "ngp_nerf_sigma_forward": (_int, [_vp, _vp, _int, _u32, _u32, _f32, _vp, _vp, _vp, _op]),
    "ngp_nerf_color_forward": (_int, [_vp, _vp, _vp, _int, _u32, _u32, _vp, _vp, _op]),
    "ngp_nerf_nets_forward": (_int, [_vp, _vp, _vp, _vp, _int, _u32, _u32, _u32, _f32, _vp, _vp, _vp, _vp, _op]),
    "ngp_nerf_color_backward": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _int, _u32, _u32, _f32, _vp, _vp, _int, _vp, _sz, _vp, _op]),
    "ngp_step_begin": (_int, [_vp, _f32, _f32, _vp]),
    "ngp_step_end": (_int, [_vp, _vp]),
    "ngp_found_inf": (_int, [_vp, _sz, _vp, _sz, _vp, _vp]),
    "ngp_adam_step": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _sz, _sz, _vp, _f32, _f32, _f32, _f32, _f32, _vp, _op]),
    "ngp_sdf_loss": (_int, [_vp, _vp, _u32, _vp, _vp, _vp, _vp, _op]),
    "ngp_sdf_forward_loss": (_int, [_vp, _vp, _int, _vp, _u32, _u32, _vp, _vp, _vp, _vp, _op]),
    "ngp_bin_points": (_int, [_vp, _vp, _u32, _f32, _u32, _vp, _sz, _vp, _vp, _vp]),
    "ngp_bin_points_workspace": (_sz, [_u32, _u32]),
    "ngp_cast_f16": (_int, [_vp, _vp, _sz, _vp]),
    "ngp_density_ema": (_int, [_vp, _vp, _f32, _u32, _vp, _vp]),
    "ngp_packbits_mean": (_int, [_vp, _u32, _vp, ctypes.c_double, _f32, _vp, _vp]),
    "ngp_umma_selftest": (_int, [_vp, _u32, _vp, _u32, _vp, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _u32, _vp]),
}


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA library has not been built (run `python torch-ngp_b200/build.py` or "
            "`python -c 'import __graft_entry__ as g; g.build()'`).  ngp_b200 has no CPU or PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


class NgpError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise NgpError(lib.ngp_last_error().decode("utf-8", "replace") or f"ngp_b200 error {rc}")


def set_option(key, value):
    check(lib.ngp_set_option(key.encode(), int(value)))


def get_option(key):
    v = _int(0)
    check(lib.ngp_get_option(key.encode(), ctypes.byref(v)))
    return v.value


def launch_count():
    return int(lib.ngp_launch_count())


def profile_collect():
    """-> {kernel name: (launch count, total ms)} since the last call (needs option profile_events=1)."""
    buf = ctypes.create_string_buffer(1 << 16)
    n = lib.ngp_profile_collect(buf, len(buf))
    if n < 0:
        check(n)
    out = {}
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.rsplit(" ", 2)
        out[name] = (int(cnt), float(ms))
    return out


def stream():
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    return None if t is None else t.data_ptr()


def dtype_code(t, what):
    if t.dtype == torch.float32:
        return DTYPE_F32
    if t.dtype == torch.float16:
        return DTYPE_F16
    if t.dtype == torch.float64:
        return DTYPE_F64
    raise RuntimeError(f"{what} must be a floating tensor")


# error wording follows the reference's TORCH_CHECK macros (gridencoder.cu:15-18, ffmlp/src/utils.h:19-23)
def check_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor")


def check_contiguous(t, name):
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be a contiguous tensor")


def check_dtype(t, dtype, name, word):
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {word} tensor")


def check_numel(t, n, name):
    if t.numel() < n:
        raise RuntimeError(f"{name} has {t.numel()} elements, expected at least {n}")


def check_tensor(t, name, dtype=None, word=None, numel=None):
    check_cuda(t, name)
    check_contiguous(t, name)
    if dtype is not None:
        check_dtype(t, dtype, name, word)
    if numel is not None:
        check_numel(t, numel, name)
