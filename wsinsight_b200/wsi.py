"""In-memory, device-resident slides exposing the reference's slide protocol.

The reference opens slides with tiffslide/openslide (``wsinsight/wsi.py:75-105``) and uses three
members: ``dimensions``, ``get_thumbnail(size)`` and ``read_region(location, level, size)``.
Decoders are out of scope here; a :class:`DeviceSlide` holds the level-0
pixels as an RGBA uint8 CUDA tensor (pitch 16-byte aligned, what the TMA-fed preprocess kernel
reads) and offers the same three members so host code written against the reference still works.
"""

from __future__ import annotations

import numpy as np

from . import _cabi


def _thumb_size(dims, size):
    """PIL ``Image.thumbnail`` aspect rule (what tiffslide.get_thumbnail ends with)."""
    import math

    w, h = dims
    x, y = size
    if x >= w and y >= h:
        return w, h

    def round_aspect(number, key):
        return max(min(math.floor(number), math.ceil(number), key=key), 1)

    aspect = w / h
    if x / y >= aspect:
        x = round_aspect(y * aspect, key=lambda n: abs(aspect - n / y))
    else:
        y = round_aspect(x / aspect, key=lambda n: 0 if n == 0 else abs(aspect - x / n))
    return x, y


class DeviceSlide:
    """Level-0 RGBA pixels resident in HBM.

    ``rgba``: CUDA uint8 tensor (H, Wp, 4) whose first ``width`` columns are valid (``Wp*4`` is
    the pitch, a multiple of 16).  ``mpp``: microns per pixel (the reference reads it from the
    file metadata, ``wsinsight/wsi.py:265-302``).
    """

    def __init__(self, rgba, width: int, height: int, mpp: float, name: str = "synthetic"):
        self.rgba = rgba
        self.width, self.height = int(width), int(height)
        self.mpp = float(mpp)
        self.name = name
        self._thumb_cache = {}

    # -- constructors ---------------------------------------------------------
    @classmethod
    def synthetic(cls, width: int, height: int, mpp: float, seed: int = 0, device=None, name=None):
        """Procedural slide generated on the device (csrc/synth.cu), never copied from the host."""
        torch = _cabi.require_cuda()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        wp = (int(width) + 3) // 4 * 4
        rgba = torch.empty((int(height), wp, 4), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            _cabi.call("wsi_synth_slide", rgba.data_ptr(), int(width), int(height), wp * 4, int(seed) & 0xFFFFFFFF,
                       _cabi.stream_ptr())
        return cls(rgba, width, height, mpp, name or f"synthetic_{width}x{height}_s{seed}")

    @classmethod
    def from_numpy(cls, arr: np.ndarray, mpp: float, device=None, name="array"):
        """(H, W, 3|4) uint8 host image -> device slide (RGB gets A=255)."""
        torch = _cabi.require_cuda()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        h, w, c = arr.shape
        wp = (w + 3) // 4 * 4
        host = np.zeros((h, wp, 4), dtype=np.uint8)
        host[:, :w, :c] = arr
        if c == 3:
            host[:, :w, 3] = 255
        return cls(torch.from_numpy(host).to(dev), w, h, mpp, name)

    # -- reference slide protocol ----------------------------------------------
    @property
    def dimensions(self):
        return (self.width, self.height)

    @property
    def pitch_bytes(self) -> int:
        return int(self.rgba.shape[1]) * 4

    def thumbnail_device(self, size=(2048, 2048), method: str = "lanczos"):
        """RGB thumbnail on the device: CUDA uint8 (Ht, Wt, 3).

        ``method="lanczos"`` (default): what the reference segments -- tiffslide ``get_thumbnail`` ->
        ``Image.thumbnail(size, LANCZOS)`` (pipeline.py:102-105), bit-exact with Pillow (csrc/thumb.cu).
        ``method="box"``: plain box average (the round-1 stand-in, kept for comparison).
        """
        import ctypes as C

        torch = _cabi.require_cuda()
        wt, ht = _thumb_size(self.dimensions, size)
        key = (wt, ht, method)
        if key not in self._thumb_cache:
            out = torch.empty((ht, wt, 3), dtype=torch.uint8, device=self.rgba.device)
            with torch.cuda.device(self.rgba.device):
                if method == "box":
                    _cabi.call("wsi_box_thumbnail", self.rgba.data_ptr(), self.width, self.height, self.pitch_bytes,
                               out.data_ptr(), ht, wt, _cabi.stream_ptr())
                elif method == "lanczos":
                    ow, oh = C.c_int(), C.c_int()
                    _cabi.call("wsi_thumbnail_geometry", self.width, self.height, int(size[0]), int(size[1]),
                               C.byref(ow), C.byref(oh), None, None)
                    assert (ow.value, oh.value) == (wt, ht), "thumbnail geometry disagrees with the Pillow aspect rule"
                    nbytes = int(_cabi.load().wsi_thumbnail_workspace_bytes(self.width, self.height, int(size[0]),
                                                                            int(size[1])))
                    ws = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=self.rgba.device)
                    _cabi.call("wsi_lanczos_thumbnail", self.rgba.data_ptr(), self.width, self.height, self.pitch_bytes,
                               int(size[0]), int(size[1]), out.data_ptr(), ws.data_ptr(), nbytes, _cabi.stream_ptr())
                    torch.cuda.current_stream().synchronize()      # the workspace is released on return
                else:
                    raise ValueError(f"unknown thumbnail method {method!r}")
            self._thumb_cache[key] = out
        return self._thumb_cache[key]

    def get_thumbnail(self, size=(2048, 2048)):
        from PIL import Image

        return Image.fromarray(self.thumbnail_device(size).cpu().numpy(), mode="RGB")

    def read_region(self, location, level, size):
        """RGBA PIL image, zero padded outside the slide (reference tests/test_all.py:747-765)."""
        from PIL import Image

        if level != 0:
            raise ValueError("DeviceSlide only holds level 0")
        x, y = int(location[0]), int(location[1])
        w, h = int(size[0]), int(size[1])
        out = np.zeros((h, w, 4), dtype=np.uint8)
        x0, y0, x1, y1 = max(x, 0), max(y, 0), min(x + w, self.width), min(y + h, self.height)
        if x1 > x0 and y1 > y0:
            out[y0 - y:y1 - y, x0 - x:x1 - x] = self.rgba[y0:y1, x0:x1].cpu().numpy()
        return Image.fromarray(out, mode="RGBA")


def synth_window_host(x0: int, y0: int, w: int, h: int, slide_w: int, slide_h: int, seed: int = 0) -> np.ndarray:
    """Regenerate a window of a synthetic slide on the HOST (bit-identical to the device generator)."""
    lib = _cabi.load()
    out = np.empty((h, w, 4), dtype=np.uint8)
    _cabi.check("wsi_synth_slide_host",
                lib.wsi_synth_slide_host(out.ctypes.data, int(x0), int(y0), int(w), int(h), int(slide_w),
                                         int(slide_h), int(seed) & 0xFFFFFFFF))
    return out
