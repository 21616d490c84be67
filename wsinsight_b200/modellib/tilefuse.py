"""Device-resident tile stitcher for the cell-model path.

Mirror of the reference's ``TileRemapStitcher`` (``wsinsight/modellib/tilefuse.py:177-415``): same constructor
arguments, the same ``accumulate_batch_torch(pred_dict, batch_coords)`` call and the same ``finalize()`` return value
(three lists: ``(1, 4)`` int32 ``[x, y, w, h]`` boxes, ``(1, K)`` float32 class means, ``(M, 2)`` int32 polygons).

What differs is where the work happens:

* the canvases ``np`` / ``hv`` / ``tp`` are a SPARSE set of ``canvas_tile`` x ``canvas_tile`` fp32 CUDA tiles allocated on
  first touch (the reference holds slide-sized host arrays and relies on the OS to keep untouched pages virtual);
  ``np_map`` / ``hv_map`` / ``tp_map`` assemble dense tensors for small slides and tests,
* ``accumulate_batch_torch`` is one fused kernel per touched canvas tile (``csrc/tilefuse.cu``): softmax, bilinear resize,
  ``hv * alpha``, ``tp`` renormalisation and the clipped paste, no host copy,
* ``finalize`` runs ``_proc_np_hv`` + ``_stitching_worker`` of every 2048-px tile (64 px of padding) on the device
  (``csrc/cellpost.cu``); only the per-instance records come back to the host.  ``release=True`` frees canvas tiles as
  soon as no later finalize job needs them.
"""

from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Tuple

import numpy as np

from .. import _cabi


class TileRemapStitcher:
    def __init__(self, n_classes: int, slide_width: int, slide_height: int, slide_patch_size: int,
                 slide_halo_size: int, slide_mpp: float, model_mpp: float, min_object_size: int = 20, device=None,
                 canvas_tile: int = 2048):
        torch = _cabi.require_cuda()
        self.n_classes = int(n_classes)
        self.slide_width, self.slide_height = int(slide_width), int(slide_height)
        self.slide_patch_size, self.slide_halo_size = int(slide_patch_size), int(slide_halo_size)
        self.alpha = model_mpp / slide_mpp
        self.min_object_size = int(min_object_size)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.type != "cuda":
            raise _cabi.WsiKernelError("TileRemapStitcher", -1, "the stitcher canvases live in HBM: a CUDA device is required")
        self.canvas_tile = int(canvas_tile)
        if self.canvas_tile < 1:
            raise ValueError("canvas_tile must be positive")
        self.tiles_x = (self.slide_width + self.canvas_tile - 1) // self.canvas_tile
        self.tiles_y = (self.slide_height + self.canvas_tile - 1) // self.canvas_tile
        # (ty, tx) -> (np (T, T), hv (T, T, 2), tp (T, T, K)) fp32 CUDA tensors, allocated on first touch
        self._tiles: Dict[Tuple[int, int], tuple] = {}
        self._staging = None

    # -- canvases ---------------------------------------------------------------------------------------------
    def _tile_dims(self, ty: int, tx: int):
        T = self.canvas_tile
        return min(T, self.slide_height - ty * T), min(T, self.slide_width - tx * T)

    def _tile(self, ty: int, tx: int):
        t = self._tiles.get((ty, tx))
        if t is None:
            torch = _cabi.require_cuda()
            h, w = self._tile_dims(ty, tx)
            t = (torch.zeros((h, w), dtype=torch.float32, device=self.device),
                 torch.zeros((h, w, 2), dtype=torch.float32, device=self.device),
                 torch.zeros((h, w, self.n_classes), dtype=torch.float32, device=self.device))
            self._tiles[(ty, tx)] = t
        return t

    def allocated_tiles(self) -> int:
        return len(self._tiles)

    def window(self, y0: int, y1: int, x0: int, x1: int):
        """Dense copies (np (h, w), hv (h, w, 2), tp (h, w, K)) of a canvas window; untouched tiles read as zero."""
        torch = _cabi.require_cuda()
        T, K = self.canvas_tile, self.n_classes
        h, w = y1 - y0, x1 - x0
        npw = torch.zeros((h, w), dtype=torch.float32, device=self.device)
        hvw = torch.zeros((h, w, 2), dtype=torch.float32, device=self.device)
        tpw = torch.zeros((h, w, K), dtype=torch.float32, device=self.device)
        for ty in range(y0 // T, (y1 - 1) // T + 1):
            for tx in range(x0 // T, (x1 - 1) // T + 1):
                t = self._tiles.get((ty, tx))
                if t is None:
                    continue
                ay0, ax0 = max(y0, ty * T), max(x0, tx * T)
                ay1, ax1 = min(y1, ty * T + t[0].shape[0]), min(x1, tx * T + t[0].shape[1])
                if ay1 <= ay0 or ax1 <= ax0:
                    continue
                src = (slice(ay0 - ty * T, ay1 - ty * T), slice(ax0 - tx * T, ax1 - tx * T))
                dst = (slice(ay0 - y0, ay1 - y0), slice(ax0 - x0, ax1 - x0))
                npw[dst] = t[0][src]
                hvw[dst] = t[1][src]
                tpw[dst] = t[2][src]
        return npw, hvw, tpw

    @property
    def np_map(self):
        return self.window(0, self.slide_height, 0, self.slide_width)[0]

    @property
    def hv_map(self):
        return self.window(0, self.slide_height, 0, self.slide_width)[1]

    @property
    def tp_map(self):
        return self.window(0, self.slide_height, 0, self.slide_width)[2]

    # -- batch step (tilefuse.py:219-278) -------------------------------------------------------------------------
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
        host = torch.as_tensor(batch_coords).detach().cpu().numpy().astype(np.int64)
        if host.ndim != 2 or host.shape[0] != b or host.shape[1] < 2:
            raise ValueError("batch_coords must be (B, >=2)")
        if b == 0:
            return
        S, T = self.slide_patch_size, self.canvas_tile
        x0 = host[:, 0] + self.slide_halo_size
        y0 = host[:, 1] + self.slide_halo_size
        # The reference pastes tile after tile, so where two interiors of one batch overlap the LATER one wins.  One launch
        # writes all tiles concurrently: keep that only when the interiors are pairwise disjoint, else paste in order.
        ox = (x0[:, None] < x0[None, :] + S) & (x0[None, :] < x0[:, None] + S)
        oy = (y0[:, None] < y0[None, :] + S) & (y0[None, :] < y0[:, None] + S)
        overlap = bool(np.triu(ox & oy, 1).any())
        groups = [np.array([i]) for i in range(b)] if overlap else [np.arange(b)]
        with torch.cuda.device(self.device):
            for g in groups:
                gx0, gy0 = x0[g], y0[g]
                for ty in range(max(0, int(gy0.min()) // T), min(self.tiles_y - 1, int(gy0.max() + S - 1) // T) + 1):
                    for tx in range(max(0, int(gx0.min()) // T), min(self.tiles_x - 1, int(gx0.max() + S - 1) // T) + 1):
                        th, tw = self._tile_dims(ty, tx)
                        hit = (gx0 < tx * T + tw) & (gx0 + S > tx * T) & (gy0 < ty * T + th) & (gy0 + S > ty * T)
                        if not hit.any():
                            continue
                        idx = g[hit]
                        lo, hi = int(idx.min()), int(idx.max()) + 1        # one launch covers the contiguous index range
                        sel = np.zeros(hi - lo, dtype=bool)
                        sel[idx - lo] = True
                        # tiles of the range that do not touch this canvas tile are moved far outside and clipped away
                        cc = np.stack([host[lo:hi, 0] - tx * T, host[lo:hi, 1] - ty * T], 1)
                        cc[~sel] = -(1 << 28)
                        coords = torch.from_numpy(cc.astype(np.int32)).to(self.device)
                        npt, hvt, tpt = self._tile(ty, tx)
                        _cabi.call("wsi_tilefuse_accumulate", f[0][lo:hi].data_ptr(), f[1][lo:hi].data_ptr(),
                                   f[2][lo:hi].data_ptr(), coords.data_ptr(), 2, hi - lo, self.n_classes, int(h), S,
                                   self.slide_halo_size, float(self.alpha), tw, th, npt.data_ptr(), hvt.data_ptr(),
                                   tpt.data_ptr(), _cabi.stream_ptr())
        # the fp32 staging copies and coordinate tensors must outlive the kernels that read them: they are kept until the
        # next call (stream order then guarantees they are no longer in use) instead of synchronising after every batch
        self._staging = (f, torch.cuda.current_stream())

    # -- finalize (tilefuse.py:280-415) ---------------------------------------------------------------------------
    @staticmethod
    def finalize_jobs(H: int, W: int, tile_size: int, padding_size: int):
        """Index-only jobs of the reference (:296-321): (pad_y0, pad_y1, pad_x0, pad_x1, interior_y0, interior_x0,
        inner_y0, inner_y1, inner_x0, inner_x1)."""
        jobs = []
        for iy0 in range(0, H, tile_size):
            for ix0 in range(0, W, tile_size):
                iy1, ix1 = min(iy0 + tile_size, H), min(ix0 + tile_size, W)
                if iy1 <= iy0 or ix1 <= ix0:
                    continue
                py0, py1 = max(0, iy0 - padding_size), min(iy1 + padding_size, H)
                px0, px1 = max(0, ix0 - padding_size), min(ix1 + padding_size, W)
                jobs.append((py0, py1, px0, px1, iy0, ix0, iy0 - py0, iy0 - py0 + (iy1 - iy0), ix0 - px0,
                             ix0 - px0 + (ix1 - ix0)))
        return jobs

    def finalize_arrays(self, tile_size: int = 2048, padding_size: int = 64, max_instances_per_tile: Optional[int] = None,
                        release: bool = False, return_inst_maps: bool = False):
        """Device finalize with packed results: ``boxes`` (N, 4) int32, ``probs`` (N, K) float32, ``poly_xy`` (P, 2) int32,
        ``poly_off`` (N + 1,) int64 (instance i owns ``poly_xy[poly_off[i]:poly_off[i + 1]]``; empty when the reference
        drops the polygon).  Tiles in raster order, instances of a tile in label order."""
        torch = _cabi.require_cuda()
        lib = _cabi.load()
        H, W, K = self.slide_height, self.slide_width, self.n_classes
        empty = (np.zeros((0, 4), np.int32), np.zeros((0, K), np.float32), np.zeros((0, 2), np.int32),
                 np.zeros(1, np.int64))
        if H <= 0 or W <= 0:
            return empty + (([],) if return_inst_maps else ())
        jobs = self.finalize_jobs(H, W, int(tile_size), int(padding_size))
        boxes, probs, polys, offs, maps = [], [], [], [np.zeros(1, np.int64)], []
        last_use: Dict[Tuple[int, int], int] = {}
        T = self.canvas_tile
        if release:
            for j, (py0, py1, px0, px1, *_r) in enumerate(jobs):
                for ty in range(py0 // T, (py1 - 1) // T + 1):
                    for tx in range(px0 // T, (px1 - 1) // T + 1):
                        last_use[(ty, tx)] = j
        with torch.cuda.device(self.device):
            for j, (py0, py1, px0, px1, iy0, ix0, a0, a1, b0, b1) in enumerate(jobs):
                th, tw = py1 - py0, px1 - px0
                touched = any((ty, tx) in self._tiles for ty in range(py0 // T, (py1 - 1) // T + 1)
                              for tx in range(px0 // T, (px1 - 1) // T + 1))
                if touched:
                    npw, hvw, tpw = self.window(py0, py1, px0, px1)
                    cap = int(max_instances_per_tile or max(1024, th * tw // 64))
                    pcap = cap * 24
                    while True:
                        ws_bytes = int(lib.wsi_cellpost_workspace_bytes(th, tw, K, cap))
                        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=self.device)
                        counts = torch.zeros(4, dtype=torch.int32, device=self.device)
                        bbox = torch.empty((cap, 4), dtype=torch.int32, device=self.device)
                        cls = torch.empty((cap, K), dtype=torch.float32, device=self.device)
                        npts = torch.empty(cap, dtype=torch.int32, device=self.device)
                        off = torch.empty(cap, dtype=torch.int32, device=self.device)
                        pxy = torch.empty((pcap, 2), dtype=torch.int32, device=self.device)
                        imap = torch.empty((th, tw), dtype=torch.int32, device=self.device) if return_inst_maps else None
                        _cabi.call("wsi_cellpost_tile", npw.data_ptr(), tw, hvw.data_ptr(), 2 * tw, tpw.data_ptr(), K * tw, K,
                                   th, tw, a0, a1, b0, b1, ix0, iy0, self.min_object_size, cap, pcap, counts.data_ptr(),
                                   bbox.data_ptr(), cls.data_ptr(), npts.data_ptr(), off.data_ptr(), pxy.data_ptr(),
                                   imap.data_ptr() if imap is not None else None, ws.data_ptr(), ws_bytes,
                                   _cabi.stream_ptr())
                        n_inst, n_pts, err, _nm = (int(v) for v in counts.cpu().tolist())
                        if err == 0 and n_inst <= cap and n_pts <= pcap:
                            break
                        cap, pcap = max(cap, 2 * n_inst), max(2 * pcap, 2 * n_pts)      # rare: grow and redo the tile
                    if n_inst:
                        boxes.append(bbox[:n_inst].cpu().numpy())
                        probs.append(cls[:n_inst].cpu().numpy())
                        polys.append(pxy[:n_pts].cpu().numpy())
                        cnt = npts[:n_inst].cpu().numpy().astype(np.int64)
                        offs.append(offs[-1][-1] + np.cumsum(cnt))
                    if return_inst_maps:
                        maps.append(imap.cpu().numpy())
                    del ws
                elif return_inst_maps:
                    maps.append(np.zeros((th, tw), np.int32))
                if release:
                    for key in [k for k, last in last_use.items() if last == j]:
                        self._tiles.pop(key, None)
                        last_use.pop(key)
        if not boxes:
            return empty + ((maps,) if return_inst_maps else ())
        out = (np.concatenate(boxes), np.concatenate(probs), np.concatenate(polys) if polys else np.zeros((0, 2), np.int32),
               np.concatenate(offs))
        return out + ((maps,) if return_inst_maps else ())

    def finalize(self, tile_size: int = 2048, padding_size: int = 64, pbar=None, num_workers: Optional[int] = None,
                 tiles_per_task: int = 1):
        """Reference signature and return value (tilefuse.py:280-415): ``(inst_all, prob_all, poly_all)`` -- lists of
        ``(1, 4)`` int32 boxes, ``(1, K)`` float32 class means and ``(M, 2)`` int32 polygons (instances whose contour has
        fewer than 3 vertices have a box and a mean but no polygon, exactly as upstream).  ``num_workers`` /
        ``tiles_per_task`` are accepted and ignored: there is no thread pool on this path."""
        boxes, probs, poly_xy, poly_off = self.finalize_arrays(tile_size, padding_size)
        inst_all: List[np.ndarray] = [b.reshape(1, -1) for b in boxes]
        prob_all: List[np.ndarray] = [p.reshape(1, -1) for p in probs]
        poly_all: List[np.ndarray] = [poly_xy[poly_off[i]:poly_off[i + 1]] for i in range(len(boxes))
                                      if poly_off[i + 1] > poly_off[i]]
        if pbar is not None:
            try:
                pbar.update(len(self.finalize_jobs(self.slide_height, self.slide_width, tile_size, padding_size)))
            except Exception:  # noqa: BLE001 - progress bars are cosmetic
                pass
        return inst_all, prob_all, poly_all
