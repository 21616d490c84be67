"""ctypes binding of libwsi_b200.so (the C ABI in include/wsi_b200.h).

There is exactly one backend.  If the shared library is missing, or no CUDA
device is present when a kernel is requested, this module raises -- there is no
CPU fallback and nothing here ever imports ``oracle``.
"""

from __future__ import annotations

import ctypes as C
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "libwsi_b200.so"

c_void_p, c_int, c_size_t, c_ll = C.c_void_p, C.c_int, C.c_size_t, C.c_longlong
c_int_p = C.POINTER(C.c_int)
c_ll_p = C.POINTER(C.c_longlong)
c_float_p = C.POINTER(C.c_float)


class WsiKernelError(RuntimeError):
    """A libwsi_b200 entry point returned a negative code."""

    def __init__(self, fn: str, code: int, msg: str):
        super().__init__(f"{fn} failed ({code}): {msg}")
        self.fn, self.code, self.msg = fn, code, msg


# name -> (restype, argtypes); mirrors include/wsi_b200.h one to one
_PROTOS = {
    "wsi_last_error": (C.c_char_p, []),
    "wsi_version": (C.c_char_p, []),
    "wsi_launch_count": (c_ll, []),
    "wsi_memcpy_d2d": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p]),
    "wsi_peer_buffer_create": (c_int, [c_size_t, C.POINTER(c_void_p), C.POINTER(C.c_ubyte)]),
    "wsi_peer_buffer_open": (c_int, [C.POINTER(C.c_ubyte), C.POINTER(c_void_p)]),
    "wsi_peer_buffer_close": (c_int, [c_void_p]),
    "wsi_peer_buffer_free": (c_int, [c_void_p]),
    "wsi_seg_workspace_bytes": (c_size_t, [c_int, c_int]),
    "wsi_seg_tissue_u8": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "wsi_seg_batch_workspace_bytes": (c_size_t, [c_int, c_int_p, c_int_p]),
    "wsi_seg_tissue_batch_u8": (c_int, [c_void_p, c_int, c_int_p, c_int_p, c_int_p, c_ll_p, c_ll_p,
                                        c_int, c_int, c_int, c_int_p, c_int_p, c_void_p, c_void_p,
                                        c_void_p, c_size_t, c_void_p]),
    "wsi_seg_tissue_batch_ex": (c_int, [c_void_p, c_int, c_int_p, c_int_p, c_int_p, c_ll_p, c_ll_p,
                                        c_int, c_int, c_int, c_int, c_int, c_int_p, c_int_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_size_t, c_void_p]),
    "wsi_outermost_u8": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "wsi_patch_grid_workspace_bytes": (c_size_t, [c_ll, c_ll, c_int, c_int]),
    "wsi_patch_grid": (c_int, [c_void_p, c_int, c_int, c_ll, c_ll, c_int, c_int, c_void_p, c_int,
                               c_void_p, c_int_p, c_void_p, c_size_t, c_void_p]),
    "wsi_patch_grid_static": (c_int, [c_void_p, c_int, c_int, c_ll, c_ll, c_int, c_int, c_void_p, c_int,
                                      c_void_p, c_int_p, c_void_p]),
    "wsi_preproc_plan_create": (c_int, [c_void_p, c_ll, c_ll, c_ll, c_int, c_int, c_float_p, c_float_p,
                                        c_int, c_int, c_int, c_int, C.POINTER(c_void_p)]),
    "wsi_preproc_run": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wsi_preproc_run_u8": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wsi_preproc_plan_info": (c_int, [c_void_p, c_int_p]),
    "wsi_preproc_plan_destroy": (None, [c_void_p]),
    "wsi_net_create": (c_int, [c_int, C.POINTER(c_void_p)]),
    "wsi_net_add_tensor": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int_p]),
    "wsi_net_add_conv": (c_int, [c_void_p, c_int, c_int, c_int, c_float_p, c_float_p, c_int, c_int, c_int,
                                 c_int, c_int, c_int, c_int, c_int, c_int]),
    "wsi_net_add_conv_slice": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_float_p, c_float_p, c_int, c_int, c_int,
                                       c_int, c_int, c_int, c_int, c_int, c_int]),
    "wsi_net_add_stem7x7": (c_int, [c_void_p, c_int, c_int, c_float_p, c_float_p, c_int, c_int]),
    "wsi_net_add_stem_parity": (c_int, [c_void_p, c_int, c_int, c_float_p, c_float_p, c_int, c_int, c_int, c_int]),
    "wsi_net_add_pool": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int]),
    "wsi_net_add_pool_bias": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                      c_float_p, c_int]),
    "wsi_net_set_conv_relu_cols": (c_int, [c_void_p, c_int]),
    "wsi_net_set_conv_residual_slice": (c_int, [c_void_p, c_int, c_int]),
    "wsi_net_set_conv_center_cols": (c_int, [c_void_p, c_int]),
    "wsi_net_add_head": (c_int, [c_void_p, c_int, c_float_p, c_float_p, c_int]),
    "wsi_net_finalize": (c_int, [c_void_p]),
    "wsi_net_forward": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "wsi_net_encode_input_map": (c_int, [c_void_p, c_int, c_void_p]),
    "wsi_net_tensor_ptr": (c_int, [c_void_p, c_int, C.POINTER(c_void_p), c_int_p]),
    "wsi_net_read_tensor": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "wsi_net_write_tensor": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "wsi_net_op_is_flat": (c_int, [c_void_p, c_int]),
    "wsi_net_num_launches": (c_int, [c_void_p]),
    "wsi_net_num_ops": (c_int, [c_void_p]),
    "wsi_net_input_geometry": (c_int, [c_void_p, c_int_p]),
    "wsi_net_num_classes": (c_int, [c_void_p]),
    "wsi_net_input_layout": (c_int, [c_void_p, c_int_p]),
    "wsi_net_input_bytes": (c_size_t, [c_void_p]),
    "wsi_preproc_plan_set_parity_layout": (c_int, [c_void_p, c_ll]),
    "wsi_preproc_plan_set_channels": (c_int, [c_void_p, c_int]),
    "wsi_preproc_plan_set_capacity": (c_int, [c_void_p, c_int]),
    "wsi_hostpipe_create": (c_int, [c_void_p, c_int, c_ll, c_int, c_float_p, c_float_p, c_int, c_int,
                                    C.POINTER(c_void_p)]),
    "wsi_hostpipe_set_host_channels": (c_int, [c_void_p, c_int]),
    "wsi_hostpipe_submit": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "wsi_hostpipe_sync": (c_int, [c_void_p]),
    "wsi_hostpipe_compute_stream": (c_void_p, [c_void_p]),
    "wsi_hostpipe_bytes": (c_int, [c_void_p, c_ll_p, c_ll_p]),
    "wsi_hostpipe_destroy": (None, [c_void_p]),
    "wsi_host_alloc_pinned": (c_int, [c_size_t, C.POINTER(c_void_p)]),
    "wsi_host_free_pinned": (c_int, [c_void_p]),
    "wsi_net_set_profiling": (c_int, [c_void_p, c_int]),
    "wsi_net_profile_read": (c_int, [c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), c_int_p, c_int,
                                     c_int_p]),
    "wsi_net_flops_per_image": (C.c_double, [c_void_p]),
    "wsi_net_conv_bytes_per_image": (C.c_double, [c_void_p]),
    "wsi_net_destroy": (None, [c_void_p]),
    "wsi_synth_slide": (c_int, [c_void_p, c_ll, c_ll, c_ll, C.c_uint, c_void_p]),
    "wsi_synth_slide_host": (c_int, [c_void_p, c_ll, c_ll, c_int, c_int, c_ll, c_ll, C.c_uint]),
    "wsi_box_thumbnail": (c_int, [c_void_p, c_ll, c_ll, c_ll, c_void_p, c_int, c_int, c_void_p]),
    "wsi_density_accumulate": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "wsi_extract_patches_rgb": (c_int, [c_void_p, c_ll, c_ll, c_ll, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "wsi_pack_patches_rgba": (c_int, [c_void_p, c_int, c_int, c_ll, c_void_p, c_void_p]),
    "wsi_pack_patches_rgba_fast": (c_int, [c_void_p, c_int, c_int, c_int, c_ll, c_void_p, c_void_p]),
    "wsi_stain_normalize_rgb": (c_int, [c_void_p, c_void_p, c_ll, C.POINTER(C.c_double), c_void_p]),
    "wsi_tilefuse_accumulate": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                        C.c_float, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "wsi_cellpost_workspace_bytes": (C.c_size_t, [c_int, c_int, c_int, c_int]),
    "wsi_cellpost_tile": (c_int, [c_void_p, c_ll, c_void_p, c_ll, c_void_p, c_ll, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_int, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "wsi_ccl_label_u8": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, C.POINTER(c_void_p), C.POINTER(c_void_p),
                                 c_void_p, c_size_t, c_void_p]),
    "wsi_thumbnail_geometry": (c_int, [c_ll, c_ll, c_int, c_int, c_int_p, c_int_p, c_int_p, c_int_p]),
    "wsi_thumbnail_workspace_bytes": (C.c_size_t, [c_ll, c_ll, c_int, c_int]),
    "wsi_lanczos_thumbnail": (c_int, [c_void_p, c_ll, c_ll, c_ll, c_int, c_int, c_void_p, c_void_p, C.c_size_t, c_void_p]),
}

EXPORTED_SYMBOLS = tuple(_PROTOS)

_lib = None


def load() -> C.CDLL:
    """dlopen libwsi_b200.so and attach prototypes.  Raises if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m wsinsight_b200.build` "
            "(there is no CPU fallback for the wsinsight_b200 kernels)"
        )
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return (load().wsi_last_error() or b"").decode("utf-8", "replace")


def check(fn_name: str, rc: int) -> None:
    if rc != 0:
        raise WsiKernelError(fn_name, rc, last_error())


def call(fn_name: str, *args):
    """Call an int-returning entry point and raise WsiKernelError on failure."""
    rc = getattr(load(), fn_name)(*args)
    check(fn_name, rc)


def require_cuda():
    """Return torch after asserting a CUDA device exists (no CPU fallback)."""
    import torch

    if not torch.cuda.is_available():
        raise RuntimeError(
            "wsinsight_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback. "
            "Set nothing: WSINFER_FORCE_CPU is not supported by this implementation."
        )
    load()
    return torch


def stream_ptr(stream=None) -> int:
    import torch

    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)


def int_array(values):
    arr = (C.c_int * len(values))(*[int(v) for v in values])
    return arr


def ll_array(values):
    arr = (C.c_longlong * len(values))(*[int(v) for v in values])
    return arr


def float_array(values):
    arr = (C.c_float * len(values))(*[float(v) for v in values])
    return arr
