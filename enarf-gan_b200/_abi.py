"""ctypes binding of libenarf_b200.so (include/enarf_b200.h).  No CPU fallback: if the library is missing
or a call fails this raises."""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libenarf_b200.so")

DTYPE_F32, DTYPE_BF16 = 0, 1
RNG_GIVEN, RNG_PHILOX = 0, 1


class Scene(ctypes.Structure):
    """Mirror of enarf_scene_t."""
    _fields_ = [
        ("B", c_int32), ("P", c_int32), ("H", c_int32), ("W", c_int32),
        ("pose", c_void_p), ("cscale", c_void_p), ("canon_pose", c_void_p),
        ("feat_planes", c_void_p), ("feat_dtype", c_int32), ("reserved0", c_int32),
        ("prob_planes", c_void_p), ("prob_batch_stride", c_int64),
        ("w1", c_void_p), ("w2", c_void_p), ("w3", c_void_p),
        ("b1", c_void_p), ("b2", c_void_p), ("b3", c_void_p),
    ]


class PairCkpt(ctypes.Structure):
    """Mirror of enarf_pair_ckpt_t."""
    _fields_ = [("rec", c_void_p), ("feat", c_void_p), ("tile_base", c_void_p), ("cursor", c_void_p),
                ("capacity", c_int64), ("skip_last", c_int32), ("reserved", c_int32), ("count", c_void_p)]


_P = c_void_p
_SIGNATURES = {
    "enarf_last_error": (c_char_p, []),
    "enarf_version": (c_int, []),
    "enarf_repack_planes": (c_int, [_P, c_int, c_int, c_int, c_int, _P, c_int, _P]),
    "enarf_unpack_plane_grads": (c_int, [_P, c_int, c_int, c_int, c_int, c_int, _P, _P]),
    "enarf_warp_planes_fwd": (c_int, [_P, _P, c_int, c_int, c_int, _P, c_int, _P]),
    "enarf_warp_planes_bwd": (c_int, [_P, c_int, _P, _P, c_int, c_int, c_int, _P, _P, _P]),
    "enarf_ray_scores": (c_int, [_P, c_int, c_int, c_int, c_int, _P, c_uint64, _P, _P, _P]),
    "enarf_ray_scores_static": (c_int, [_P, c_int, c_int, c_int, c_int, _P, c_uint64, _P, _P]),
    "enarf_select_rays": (c_int, [_P, c_int, c_int, c_int, c_int, _P, _P, _P]),
    "enarf_photometric_loss": (c_int, [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, c_float, _P, _P, _P, _P]),
    "enarf_near_far": (c_int, [_P, c_int, c_int, c_float, c_float, _P, _P]),
    "enarf_depth_range": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, _P, _P, _P, _P, _P]),
    "enarf_render_coarse": (c_int, [POINTER(Scene), _P, _P, _P, c_int, c_int, _P, _P, _P, c_float, _P, _P, _P]),
    "enarf_fine_sample": (c_int, [_P, c_int, c_int, c_int, c_int, _P, _P, c_uint64, _P, c_int64, _P, _P, _P]),
    "enarf_fine_tiles_per_sample": (c_int, [c_int, c_int]),
    "enarf_render_fine_fwd": (c_int, [POINTER(Scene), _P, _P, _P, _P, c_int, c_int, _P, _P, _P, _P, _P, _P, _P,
                                      POINTER(PairCkpt), c_float, _P, _P, c_float, _P]),
    "enarf_render_fine_bwd": (c_int, [POINTER(Scene), _P, _P, _P, _P, _P, _P, POINTER(PairCkpt), c_int, c_int, _P, _P, _P, _P,
                                      _P, _P, c_int64, _P, _P, _P, _P, _P, _P, c_float, c_float, _P]),
    "enarf_density_query": (c_int, [POINTER(Scene), _P, c_int, _P, _P, _P, _P]),
    "enarf_triplane_sample_fwd": (c_int, [_P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                          _P, _P]),
    "enarf_triplane_sample_bwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                          c_int, _P, _P, _P]),
    "enarf_triplane_sample_cl_fwd": (c_int, [_P, _P, c_int, c_int, c_int, c_int64, c_int, c_int, _P, _P]),
    "enarf_triplane_sample_cl_bwd": (c_int, [_P, _P, _P, c_int, c_int, c_int, c_int64, c_int, c_int, _P, _P, _P]),
}
EXPORTS = tuple(_SIGNATURES)

_lib = None


def lib():
    """The loaded library (loads on first use)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  There is no CPU fallback.")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().enarf_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libenarf_b200 error {rc}: {msg or what}")


def ptr(t, dtype=torch.float32, allow_none=False):
    """Device pointer of a contiguous CUDA tensor (None -> NULL when allowed)."""
    if t is None:
        if allow_none:
            return None
        raise ValueError("tensor required")
    if not t.is_cuda:
        raise ValueError("libenarf_b200 needs CUDA tensors (there is no CPU path)")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"expected {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError("tensor must be contiguous")
    if t.device.index != torch.cuda.current_device():
        # the library launches on the current device's stream: refuse rather than launch on the wrong GPU
        raise RuntimeError(f"tensor on {t.device} but the current device is cuda:{torch.cuda.current_device()}: call "
                           "under `with torch.cuda.device(tensor.device)` (rendering.render / render_rays do)")
    return t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream


def on_device_of(index_or_name):
    """Decorator: run the wrapped function with the CUDA device of one of its tensor arguments current, so that a
    caller who never called torch.cuda.set_device still launches on the GPU that owns the tensors."""
    import functools

    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*args, **kwargs):
            t = kwargs.get(index_or_name) if isinstance(index_or_name, str) else args[index_or_name]
            if isinstance(t, torch.Tensor) and t.is_cuda and t.device.index != torch.cuda.current_device():
                with torch.cuda.device(t.device):
                    return fn(*args, **kwargs)
            return fn(*args, **kwargs)
        return wrapper
    return deco
