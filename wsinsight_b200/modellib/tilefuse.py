"""Device-resident tile stitcher for the cell-model path.

Mirror of the batch step of the reference's ``TileRemapStitcher`` (``wsinsight/modellib/tilefuse.py:177-278``):
same constructor arguments, same ``accumulate_batch_torch(pred_dict, batch_coords)`` call and the same canvases
``np_map`` (H, W), ``hv_map`` (H, W, 2), ``tp_map`` (H, W, K) -- but the canvases are CUDA tensors written by one
fused kernel (``csrc/tilefuse.cu``) instead of host arrays filled by a D2H copy and a Python paste loop.

``finalize`` (per-tile watershed, instance bounding boxes and contours, ``tilefuse.py:280-415``) is CPU
post-processing of the cell workflow and is out of scope: it raises ``NotImplementedError``.
"""

from __future__ import annotations

from .. import _cabi


class TileRemapStitcher:
    def __init__(self, n_classes: int, slide_width: int, slide_height: int, slide_patch_size: int,
                 slide_halo_size: int, slide_mpp: float, model_mpp: float, min_object_size: int = 20, device=None):
        torch = _cabi.require_cuda()
        self.n_classes = int(n_classes)
        self.slide_width, self.slide_height = int(slide_width), int(slide_height)
        self.slide_patch_size, self.slide_halo_size = int(slide_patch_size), int(slide_halo_size)
        self.alpha = model_mpp / slide_mpp
        self.min_object_size = int(min_object_size)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.type != "cuda":
            raise _cabi.WsiKernelError("TileRemapStitcher", -1, "the stitcher canvases live in HBM: a CUDA device is required")
        h, w, k = self.slide_height, self.slide_width, self.n_classes
        self.np_map = torch.zeros((h, w), dtype=torch.float32, device=self.device)
        self.hv_map = torch.zeros((h, w, 2), dtype=torch.float32, device=self.device)
        self.tp_map = torch.zeros((h, w, k), dtype=torch.float32, device=self.device)

    def accumulate_batch_torch(self, pred_dict: dict, batch_coords):
        """pred_dict: {"np": (B,2,h,h), "hv": (B,2,h,h), "tp": (B,K,h,h)} CUDA tensors (or the reference's
        alternative keys); batch_coords: (B, >=2) [minx, miny, ...] of each tile in canvas pixels."""
        torch = _cabi.require_cuda()
        if "np" in pred_dict and "hv" in pred_dict and "tp" in pred_dict:
            np_logits, hv, tp_logits = pred_dict["np"], pred_dict["hv"], pred_dict["tp"]
        elif "nuclei_binary_map" in pred_dict and "hv_map" in pred_dict and "nuclei_type_map" in pred_dict:
            np_logits, hv, tp_logits = pred_dict["nuclei_binary_map"], pred_dict["hv_map"], pred_dict["nuclei_type_map"]
        else:
            raise AssertionError("pred_dict needs np/hv/tp (or nuclei_binary_map/hv_map/nuclei_type_map)")
        b, _, h, w = np_logits.shape
        if w != h or tuple(hv.shape) != (b, 2, h, h) or tuple(tp_logits.shape) != (b, self.n_classes, h, h) \
                or np_logits.shape[1] != 2:
            raise ValueError("expected np (B,2,h,h), hv (B,2,h,h), tp (B,K,h,h)")
        f = [t.to(self.device, torch.float32).contiguous() for t in (np_logits, hv, tp_logits)]
        coords = torch.as_tensor(batch_coords).to(self.device).to(torch.int32).contiguous()
        if coords.dim() != 2 or coords.shape[0] != b or coords.shape[1] < 2:
            raise ValueError("batch_coords must be (B, >=2)")
        with torch.cuda.device(self.device):
            _cabi.call("wsi_tilefuse_accumulate", f[0].data_ptr(), f[1].data_ptr(), f[2].data_ptr(), coords.data_ptr(),
                       int(coords.shape[1]), int(b), self.n_classes, int(h), self.slide_patch_size, self.slide_halo_size,
                       float(self.alpha), self.slide_width, self.slide_height, self.np_map.data_ptr(),
                       self.hv_map.data_ptr(), self.tp_map.data_ptr(), _cabi.stream_ptr())
        torch.cuda.current_stream().synchronize()      # the fp32 / int32 staging copies are released on return

    def finalize(self, *args, **kwargs):
        raise NotImplementedError("watershed instance extraction (tilefuse.py:280-415) belongs to the cell workflow, "
                                  "which is outside the patch-classification path")
