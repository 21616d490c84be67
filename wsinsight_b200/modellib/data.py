"""Patch data path on the B200: coords -> bf16 NHWC batches, one fused kernel per batch.

Mirror of ``wsinsight/modellib/data.py``: ``_read_patch_coords`` (data.py:22-63) and
``WholeSlideImagePatches`` (data.py:149-322).  The reference's ``__getitem__`` does
``read_region`` -> ``convert("RGB")`` -> torchvision transform per patch in DataLoader workers;
here the slide already lives in HBM (:class:`wsinsight_b200.wsi.DeviceSlide`) and
``wsi_preproc_run`` (csrc/preproc.cu) produces a whole batch per launch.
"""

from __future__ import annotations

import ctypes as C

import numpy as np

from .. import _cabi
from .transforms import FusedTransform


def _read_patch_coords(path) -> np.ndarray:
    """(N, 4) int [minx, miny, width, height] from a patch file (data.py:22-63).

    Reads the reference's HDF5 layout when ``h5py`` is importable, else the ``.npz`` mirror
    written by :func:`wsinsight_b200.patchlib.io.save_hdf5`.
    """
    from ..patchlib.io import load_patch_file

    d = load_patch_file(path)
    coords = np.asarray(d["coords"])
    if coords.ndim != 2:
        raise ValueError(f"expected coords to have 2 dimensions, got {coords.ndim}")
    if coords.shape[1] != 2:
        raise ValueError(f"expected second dim of coords to have len 2 but got {coords.shape[1]}")
    if int(d["patch_level"]) != 0:
        raise NotImplementedError(f"This script is designed for patch_level=0 but got {d['patch_level']}")
    ps = int(d["patch_size"])
    wh = np.full_like(coords, ps)
    return np.concatenate((coords, wh), axis=1)


class PreprocessPlan:
    """Owns a ``wsi_preproc_plan`` for one (slide, patch_size, transform, output layout)."""

    def __init__(self, slide, patch_size: int, transform: FusedTransform, out_H=None, out_W=None, pad_top=0,
                 pad_left=0, plane_stride_px=0, channels=4):
        torch = _cabi.require_cuda()
        self.slide = slide
        self.patch_size = int(patch_size)
        self.S = transform.out_size(patch_size)
        self.out_H = int(out_H) if out_H is not None else self.S
        self.out_W = int(out_W) if out_W is not None else self.S
        self.pad_top, self.pad_left = int(pad_top), int(pad_left)
        self._capacity = 0
        self._h = C.c_void_p()
        mean = _cabi.float_array(transform.mean) if transform.mean is not None else None
        std = _cabi.float_array(transform.std) if transform.std is not None else None
        with torch.cuda.device(slide.rgba.device):
            _cabi.call("wsi_preproc_plan_create", slide.rgba.data_ptr(), slide.width, slide.height,
                       slide.pitch_bytes, self.patch_size, self.S, mean, std, self.out_H, self.out_W, self.pad_top,
                       self.pad_left, C.byref(self._h))
            self.plane_stride_px = int(plane_stride_px)
            if self.plane_stride_px:
                _cabi.call("wsi_preproc_plan_set_parity_layout", self._h, self.plane_stride_px)
            self.channels = int(channels)
            if self.channels != 4:
                _cabi.call("wsi_preproc_plan_set_channels", self._h, self.channels)

    def info(self):
        a = (C.c_int * 8)()
        _cabi.call("wsi_preproc_plan_info", self._h, a)
        keys = ("TR", "nbands", "G", "ngroups", "box_w", "box_h", "max_src_rows", "smem_bytes")
        return dict(zip(keys, [int(v) for v in a]))

    def run(self, coords_xy, out, stream=None):
        """coords_xy: CUDA int32 (n, 2) top-left corners; out: bf16 buffer, [>=n][out_H][out_W][4] or, with
        the parity layout, the flat buffer of ``B200Network.alloc_input``."""
        torch = _cabi.require_cuda()
        n = int(coords_xy.shape[0])
        assert coords_xy.dtype == torch.int32 and coords_xy.is_contiguous() and coords_xy.is_cuda
        assert out.dtype == torch.bfloat16 and out.is_contiguous()
        if self.plane_stride_px:
            assert out.numel() >= 2 * self.plane_stride_px * 4
        else:
            cap = int(out.numel() // (self.out_H * self.out_W * self.channels))
            if cap != self._capacity:
                _cabi.call("wsi_preproc_plan_set_capacity", self._h, cap)
                self._capacity = cap
        with torch.cuda.device(out.device):
            _cabi.call("wsi_preproc_run", self._h, coords_xy.data_ptr(), n, out.data_ptr(), _cabi.stream_ptr(stream))

    def run_u8(self, coords_xy, stream=None):
        """Parity helper: the resized uint8 patches (n, S, S, 4) with channel 3 = 0."""
        torch = _cabi.require_cuda()
        n = int(coords_xy.shape[0])
        out = torch.empty((n, self.S, self.S, 4), dtype=torch.uint8, device=coords_xy.device)
        with torch.cuda.device(out.device):
            _cabi.call("wsi_preproc_run_u8", self._h, coords_xy.data_ptr(), n, out.data_ptr(),
                       _cabi.stream_ptr(stream))
        return out

    def close(self):
        if self._h:
            _cabi.load().wsi_preproc_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class WholeSlideImagePatches:
    """Batch iterator over the patches of one slide (reference class name, data.py:149-322).

    Iterating yields ``(batch_bf16, batch_coords)``: the device batch in the network's input
    layout and the matching ``(b, 4)`` int array ``[minx, miny, width, height]``.
    """

    def __init__(self, slide, coords_xywh: np.ndarray, transform: FusedTransform, batch_size: int,
                 input_layout=None, device_coords=None):
        torch = _cabi.require_cuda()
        self.slide = slide
        self.coords = np.asarray(coords_xywh)
        if self.coords.ndim != 2 or self.coords.shape[1] != 4:
            raise ValueError("coords must be (N, 4) [minx, miny, width, height]")
        ps = np.unique(self.coords[:, 2:]) if len(self.coords) else np.array([0])
        if len(ps) != 1:
            raise ValueError("all patches of a slide must share one square patch size")
        self.patch_size = int(ps[0])
        self.batch_size = int(batch_size)
        lay = input_layout or {}
        self.plan = PreprocessPlan(slide, self.patch_size, transform, lay.get("out_H"), lay.get("out_W"),
                                   lay.get("pad_top", 0), lay.get("pad_left", 0), lay.get("plane_stride_px", 0),
                                   lay.get("channels", 4))
        dev = slide.rgba.device
        self.dev_coords = device_coords if device_coords is not None else \
            torch.from_numpy(np.ascontiguousarray(self.coords[:, :2], dtype=np.int32)).to(dev)
        nelem = lay.get("buffer_elems") or self.batch_size * self.plan.out_H * self.plan.out_W * self.plan.channels
        self._buf = torch.zeros(int(nelem), dtype=torch.bfloat16, device=dev)

    def __len__(self):
        return (len(self.coords) + self.batch_size - 1) // self.batch_size

    def __iter__(self):
        for s in range(0, len(self.coords), self.batch_size):
            e = min(s + self.batch_size, len(self.coords))
            self.plan.run(self.dev_coords[s:e], self._buf)
            yield self._buf, self.coords[s:e]
