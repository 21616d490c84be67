"""Patch classification over precomputed patch files (classification branch only).

Mirror of ``wsinsight/modellib/run_inference.py``: signature :45-65, per-slide loop :205-225,
classification loop :537-560, CSV assembly :565-607,732-733.  The reference moves fp32 NCHW
batches host->device, runs the TorchScript module through cuDNN and does softmax on the CPU;
here each batch is two library calls on one stream (fused preprocess, network forward with the
softmax in the tail kernel) and one device->host copy of the probabilities per slide.

Multi-GPU: the reference wraps the model in ``nn.DataParallel`` (:173-174).  Here ranks of a
``torch.distributed`` job each take a contiguous range of the x-major sorted patch list (an
x-band of the slide) and rank 0 gathers the ``(N_r, K)`` probabilities with one collective.
"""

from __future__ import annotations

import logging
from pathlib import Path
from typing import Dict, List, Optional

import numpy as np

from .. import _cabi
from ..errors import PatchDirectoryNotFound, ResultsDirectoryNotFound
from .data import WholeSlideImagePatches, _read_patch_coords
from .models import Model, get_pretrained_torch_module
from .transforms import make_compose_from_transform_config

logger = logging.getLogger(__name__)


def shard_range(n: int, rank: int, world: int):
    """Contiguous range [lo, hi) of rank ``rank``: N*r/R .. N*(r+1)/R."""
    return (n * rank) // world, (n * (rank + 1)) // world


def classify_patches_device(slide, coords_xywh: np.ndarray, net, transform, batch_size: int, device_coords=None,
                            probs_out=None):
    """Hot loop for one slide on one GPU: returns a CUDA fp32 (N, K) tensor of probabilities.

    Everything is enqueued on the current stream; nothing synchronises until the caller reads
    the result.  (reference: run_inference.py:537-560)
    """
    torch = _cabi.require_cuda()
    n = len(coords_xywh)
    K = net.num_classes
    probs = probs_out if probs_out is not None else torch.empty((n, K), dtype=torch.float32, device=net.device)
    if n == 0:
        return probs
    # the network's input buffer holds at most max_batch patches (the parity-plane layout fixes the plane stride)
    batch_size = max(1, min(int(batch_size), net.max_batch))
    dset = WholeSlideImagePatches(slide, coords_xywh, transform, batch_size, net.input_layout, device_coords)
    s = 0
    for batch, bc in dset:
        b = len(bc)
        net.forward_into(batch, probs[s:s + b])
        s += b
    return probs


def classify_patches_from_images(images, net, transform, batch_size: int, probs_out=None):
    """Classification from the cached ``/images`` patches instead of the slide (reference ``data.py:243-268,293-300``:
    ``_get_patch_from_hdf5`` -> transform -> model).  ``images``: uint8 (N, ps, ps, 3), CUDA tensor or numpy array.

    Each batch is packed into an RGBA stack on the device (``wsi_pack_patches_rgba``) that the fused preprocess
    kernel reads as a slide of width ``ps`` with patch *i* at ``(0, i * ps)``; the result is bit-identical to
    :func:`classify_patches_device` on the same pixels."""
    from ..wsi import DeviceSlide
    from .data import PreprocessPlan

    torch = _cabi.require_cuda()
    if not hasattr(images, "is_cuda"):
        images = torch.from_numpy(np.ascontiguousarray(images))
    images = images.to(net.device).contiguous()
    if images.dim() != 4 or images.shape[3] != 3 or images.shape[1] != images.shape[2] or images.dtype != torch.uint8:
        raise ValueError(f"images must be uint8 (N, ps, ps, 3), got {tuple(images.shape)} {images.dtype}")
    n, ps = int(images.shape[0]), int(images.shape[1])
    K = net.num_classes
    probs = probs_out if probs_out is not None else torch.empty((n, K), dtype=torch.float32, device=net.device)
    if n == 0:
        return probs
    B = min(int(batch_size), net.max_batch)
    wp = (ps + 3) // 4 * 4
    stack = torch.zeros((B * ps, wp, 4), dtype=torch.uint8, device=net.device)
    view = DeviceSlide(stack, ps, B * ps, mpp=1.0, name="image-cache")
    lay = net.input_layout
    plan = PreprocessPlan(view, ps, transform, lay["out_H"], lay["out_W"], lay["pad_top"], lay["pad_left"],
                          lay.get("plane_stride_px", 0), lay.get("channels", 4))
    coords = torch.zeros((B, 2), dtype=torch.int32, device=net.device)
    coords[:, 1] = torch.arange(B, dtype=torch.int32, device=net.device) * ps
    xin = net.alloc_input()
    with torch.cuda.device(net.device):
        for s in range(0, n, B):
            b = min(B, n - s)
            _cabi.call("wsi_pack_patches_rgba", images[s:s + b].data_ptr(), b, ps, wp * 4, stack.data_ptr(),
                       _cabi.stream_ptr())
            plan.run(coords[:b], xin)
            net.forward_into(xin, probs[s:s + b])
    torch.cuda.current_stream().synchronize()      # stack / plan / xin are released on return
    plan.close()
    return probs


def classify_patches_stain_normalized(slide, coords_xywh: np.ndarray, net, transform, batch_size: int, W_est, W_def,
                                      chunk: int = 4096):
    """Classification with Macenko stain normalisation (reference ``data.py:293-300``): patches are gathered from the
    slide, normalised per pixel on the device (``wsi_stain_normalize_rgb``) and classified from the patch stack, chunk
    by chunk.  Returns a CUDA fp32 (N, K) tensor."""
    from ..patchlib.io import extract_patches_from_slide
    from .stain import normalize_patches_device

    torch = _cabi.require_cuda()
    n = len(coords_xywh)
    probs = torch.empty((n, net.num_classes), dtype=torch.float32, device=net.device)
    if n == 0:
        return probs
    ps = int(coords_xywh[0, 2])
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        px = extract_patches_from_slide(slide, coords_xywh[s:e, :2], ps)
        normalize_patches_device(px, W_est, W_def, out=px)
        classify_patches_from_images(px, net, transform, batch_size, probs_out=probs[s:e])
    return probs


def gather_probs(probs_local, n_total: int, rank: int, world: int):
    """Gather per-rank (N_r, K) probability shards on rank 0 (NCCL on GPUs, gloo on CPU).

    Shards follow :func:`shard_range`, so ranks may differ by one row: shards are padded to the
    largest, all-gathered, and trimmed.  Returns the (N, K) tensor on rank 0, ``None`` elsewhere.
    """
    import torch
    import torch.distributed as dist

    if world == 1:
        return probs_local
    K = int(probs_local.shape[1])
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((mx, K), dtype=probs_local.dtype, device=probs_local.device)
    pad[: probs_local.shape[0]] = probs_local
    bufs = [torch.empty_like(pad) for _ in range(world)] if rank == 0 else None
    dist.gather(pad, bufs, dst=0)
    if rank != 0:
        return None
    return torch.cat([bufs[r][: hi - lo] for r, (lo, hi) in enumerate(sizes)], dim=0)


def dist_rank_world():
    """(rank, world) of the running ``torch.distributed`` job, (0, 1) outside one."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class PeerResultBuffer:
    """The (N, K) fp32 probability array of one slide, owned by rank 0 and mapped into every rank of the node through
    CUDA IPC (``wsi_peer_buffer_*``): each rank's tail kernel (GAP + FC + softmax) stores its rows straight into rank 0's
    memory over NVLink, so the result path needs no gather collective -- only a barrier before rank 0 reads."""

    _cache = {}      # (K, world, device index) -> PeerResultBuffer kept mapped between slides

    @classmethod
    def get(cls, n: int, K: int, rank: int, world: int, device):
        """A buffer with room for ``n`` rows: the mapped buffer of the previous slide when it is large enough (every rank
        sees the same ``n``, so every rank takes the same branch), else a new one of twice the size."""
        torch = _cabi.require_cuda()
        key = (int(K), int(world), torch.device(device).index)
        buf = cls._cache.get(key)
        if buf is None or buf.capacity < n:
            if buf is not None:
                buf.release()
            buf = cls(max(2 * int(n), 1 << 16), K, rank, world, device)
            cls._cache[key] = buf
        buf.n = int(n)
        return buf

    def __init__(self, n: int, K: int, rank: int, world: int, device):
        import ctypes as C

        import torch.distributed as dist

        torch = _cabi.require_cuda()
        self.rank, self.world, self.n, self.K = rank, world, int(n), int(K)
        self.capacity = int(n)
        self._ptr = C.c_void_p()
        self.tensor = None
        handle = (C.c_ubyte * 64)()
        nbytes = max(1, self.n * self.K) * 4
        err = None
        with torch.cuda.device(device):
            # every step that can fail on ONE rank is followed by a collective vote, so that either all ranks use the
            # buffer or all ranks raise (a one-sided failure would leave the others waiting in a barrier)
            if rank == 0:
                try:
                    _cabi.call("wsi_peer_buffer_create", nbytes, C.byref(self._ptr), handle)
                except Exception as e:  # noqa: BLE001
                    err = e
            box = [bytes(handle)]
            dist.broadcast_object_list(box, src=0)
            if rank != 0 and err is None:
                try:
                    h = (C.c_ubyte * 64).from_buffer_copy(box[0])
                    _cabi.call("wsi_peer_buffer_open", h, C.byref(self._ptr))
                except Exception as e:  # noqa: BLE001
                    err = e
            if err is None:
                try:
                    class _Raw:      # zero-copy tensor view of the mapped memory
                        pass

                    raw = _Raw()
                    raw.__cuda_array_interface__ = {"shape": (self.n, self.K), "typestr": "<f4",
                                                    "data": (int(self._ptr.value), False), "version": 2}
                    self.tensor = torch.as_tensor(raw)       # the device is read off the pointer; must be zero-copy
                    if self.n and int(self.tensor.data_ptr()) != int(self._ptr.value):
                        raise RuntimeError("the tensor view of the mapped buffer is not zero-copy")
                except Exception as e:  # noqa: BLE001
                    err = e
            bad = torch.tensor([0 if err is None else 1], dtype=torch.int32, device=torch.device(device))
            dist.all_reduce(bad, op=dist.ReduceOp.MAX)
            if int(bad.item()):
                self.tensor = None
                if self._ptr.value:
                    try:
                        _cabi.call("wsi_peer_buffer_close" if rank != 0 else "wsi_peer_buffer_free", self._ptr)
                    except Exception:  # noqa: BLE001
                        pass
                raise RuntimeError(f"peer result buffer unavailable on this node (rank {rank}: {err})")

    def rows(self, lo: int, hi: int):
        return self.tensor[lo:hi]

    def collect(self):
        """Barrier, then the first ``n`` rows as a NumPy array on rank 0 (None elsewhere).  The buffer stays mapped for the
        next slide; a second barrier keeps the other ranks from overwriting it before rank 0 has copied."""
        import torch.distributed as dist

        torch = _cabi.require_cuda()
        torch.cuda.synchronize()
        dist.barrier()                     # every rank's stores have landed
        out = self.tensor[: self.n].cpu().numpy().copy() if self.rank == 0 else None
        dist.barrier()                     # rank 0 has its copy: the rows may be rewritten
        return out

    def release(self):
        """Unmap (ranks > 0) / free (rank 0) the buffer; collective."""
        import torch.distributed as dist

        self.tensor = None
        if self.rank != 0 and self._ptr.value:
            _cabi.call("wsi_peer_buffer_close", self._ptr)
        dist.barrier()                     # all mappings closed before the owner frees
        if self.rank == 0 and self._ptr.value:
            _cabi.call("wsi_peer_buffer_free", self._ptr)
        self._ptr.value = None


def _result_path(result_path=None) -> str:
    import os

    rp = (result_path or os.environ.get("WSI_B200_RESULT_PATH") or "nccl").lower()
    if rp not in ("nccl", "peer"):
        raise ValueError(f"result path must be 'nccl' or 'peer', got {rp!r}")
    return rp


def classify_coords_sharded(slide, coords_xywh: np.ndarray, net, transform, batch_size: int, rank=None, world=None,
                            result_path=None):
    """ONE slide on all ranks of a ``torch.distributed`` job (replaces ``nn.DataParallel``, reference
    run_inference.py:173-174): rank r classifies ``shard_range(N, r, R)`` of the x-major patch list -- an x-band of
    the slide, no data-path collective -- and the ``(N_r, K)`` shards are gathered on rank 0 with one collective.

    ``result_path="peer"`` (or ``WSI_B200_RESULT_PATH=peer``): no collective at all -- the probabilities are stored into
    a rank-0 buffer mapped into every rank (:class:`PeerResultBuffer`); default ``"nccl"``.

    Returns the (N, K) float32 NumPy array on rank 0, ``None`` on the other ranks."""
    if rank is None or world is None:
        rank, world = dist_rank_world()
    n = len(coords_xywh)
    lo, hi = shard_range(n, rank, world)
    if world > 1 and _result_path(result_path) == "peer":
        buf = PeerResultBuffer.get(n, net.num_classes, rank, world, net.device)
        classify_patches_device(slide, coords_xywh[lo:hi], net, transform, batch_size, probs_out=buf.rows(lo, hi))
        return buf.collect()
    local = classify_patches_device(slide, coords_xywh[lo:hi], net, transform, batch_size)
    full = gather_probs(local, n, rank, world)
    return full.cpu().numpy() if full is not None else None


def classify_slide_sharded(slide, patch_size_px: int, patch_spacing_um_px: float, net, transform, batch_size: int,
                           rank=None, world=None, thumb=None, result_path=None, **seg_kwargs):
    """Whole-slide form: tissue segmentation + patchify run redundantly on every rank (about a millisecond; the
    coordinates then need no broadcast), then :func:`classify_coords_sharded`.

    Returns ``(coords_xywh (N, 4) int64, probs (N, K) float32)`` on rank 0, ``(coords_xywh, None)`` elsewhere."""
    from ..patchlib import segment_and_patch_slides_device

    res = segment_and_patch_slides_device([slide], patch_size_px, patch_spacing_um_px,
                                          thumbs=None if thumb is None else [thumb], **seg_kwargs)[0]
    if res.coords is None:
        return np.zeros((0, 4), np.int64), (np.zeros((0, net.num_classes), np.float32) if (rank or 0) == 0 else None)
    if rank is None or world is None:
        rank, world = dist_rank_world()
    n = int(res.coords.shape[0])
    lo, hi = shard_range(n, rank, world)
    cxy = res.coords.cpu().numpy().astype(np.int64)
    xywh = np.concatenate([cxy, np.full_like(cxy, res.patch_size)], 1)
    dc = res.coords[lo:hi].contiguous()
    if world > 1 and _result_path(result_path) == "peer":
        buf = PeerResultBuffer.get(n, net.num_classes, rank, world, net.device)
        classify_patches_device(slide, xywh[lo:hi], net, transform, batch_size, device_coords=dc,
                                probs_out=buf.rows(lo, hi))
        return xywh, buf.collect()
    local = classify_patches_device(slide, xywh[lo:hi], net, transform, batch_size, device_coords=dc)
    full = gather_probs(local, n, rank, world)
    return xywh, (full.cpu().numpy() if full is not None else None)


def write_slide_csv(path, coords_xywh: np.ndarray, probs: np.ndarray, class_names: List[str]) -> None:
    """CSV assembly of run_inference.py:568-607,732-733: minx,miny,width,height,prob_<class>..."""
    import pandas as pd

    df = pd.DataFrame(dict(minx=coords_xywh[:, 0], miny=coords_xywh[:, 1], width=coords_xywh[:, 2],
                           height=coords_xywh[:, 3]))
    probs = np.asarray(probs)
    if probs.ndim == 1:
        probs = probs[:, None]
    for i, c in enumerate(class_names):
        df[f"prob_{c}"] = probs[:, i]
    Path(path).parent.mkdir(parents=True, exist_ok=True)
    df.to_csv(path, index=False)


def run_inference(
    wsi_dir,
    slide_paths,
    results_dir,
    references_dir=None,
    qupath_detection_dir=None,
    qupath_geojson_detection_dir=None,
    qupath_geojson_annotation_dir=None,
    qupath_name_as_class: bool = False,
    model_info: Model = None,
    halo_size_px: int = 46,
    batch_size: int = 32,
    num_workers: int = 4,
    speedup: bool = False,
    stain_normalization: bool = False,
    object_based: bool = False,
    object_detection: Optional[str] = None,
    mixed_precision: bool = False,
    stitch_workers: Optional[int] = None,
    slides: Optional[Dict[str, object]] = None,
):
    """Reference signature (run_inference.py:45-65).  ``slides`` maps slide path -> opened
    :class:`~wsinsight_b200.wsi.DeviceSlide` (decoders are out of scope).  ``num_workers`` is
    accepted and ignored: there are no DataLoader workers on this path.

    Under ``torch.distributed`` (one process per GPU, NCCL) every rank calls this function with the same arguments
    and its own ``slides`` (each slide open on the rank's device): each slide's patch list is sharded across the
    ranks (:func:`classify_coords_sharded`) and rank 0 writes the CSV -- the reference's ``nn.DataParallel``
    (:173-174) without the scatter/gather of image batches.

    Returns ``(failed_patching, failed_inference)`` like the reference.
    """
    if references_dir or qupath_detection_dir or qupath_geojson_detection_dir or qupath_geojson_annotation_dir \
            or object_based or object_detection:
        raise NotImplementedError("QuPath / reference-annotation / object-based branches are outside the B200 hot path")
    from ..patchlib.io import load_patch_file

    results_dir = Path(results_dir)
    if not results_dir.exists():
        raise ResultsDirectoryNotFound(results_dir)
    patch_dir = results_dir / "patches"
    if not patch_dir.exists():
        raise PatchDirectoryNotFound("Results dir must include 'patches' dir")
    model_output_dir = results_dir / "model-outputs-csv"
    model_output_dir.mkdir(exist_ok=True)
    rank, world = dist_rank_world()
    slides = slides or {}
    patch_paths = sorted(list(patch_dir.glob("*.h5")) + list(patch_dir.glob("*.h5.npz")))
    failed_patching = [str(p) for p in (slide_paths or [])
                       if not any(Path(str(q).replace(".npz", "")).stem == Path(str(p)).stem for q in patch_paths)]
    failed_inference: List[str] = []
    transform = make_compose_from_transform_config(model_info.config.transform)
    net = None
    for patch_path in patch_paths:
        # the reference isolates failures per slide and records ``wsi_path.stem`` (run_inference.py:232-280)
        stem = Path(str(patch_path).replace(".npz", "")).stem
        try:
            meta = load_patch_file(patch_path)
            wsi_path = str(meta["slide_path"])
            stem = Path(wsi_path).stem
        except Exception as e:  # noqa: BLE001 - one unreadable patch file must not abort the cohort
            failed_inference.append(stem)
            logger.error(f"Cannot read patch file {patch_path}", exc_info=e)
            continue
        slide_csv = model_output_dir / (stem + ".csv")
        if slide_csv.exists():
            logger.info(f"Output CSV exists... skipping. {slide_csv}")
            continue
        slide = slides.get(wsi_path)
        if slide is None:
            failed_inference.append(stem)
            logger.error(f"slide {wsi_path} is not open on the device")
            continue
        try:
            coords = _read_patch_coords(patch_path)
            if len(coords) == 0:
                continue
            if net is None:
                S = transform.out_size(int(coords[0, 2]))
                from .engine import B200Network

                net = B200Network.from_torchscript(model_info.model_path, model_info.config.num_classes,
                                                   input_size=S, max_batch=batch_size, device=slide.rgba.device)
            if stain_normalization:
                # run_inference.py:232-264: one stain matrix per slide from a 256-patch sample, eosin / hematoxylin target
                from .stain import default_target_matrix, estimate_slide_stain_matrix

                W_est = estimate_slide_stain_matrix(slide, coords, 256, seed=0)
                lo, hi = shard_range(len(coords), rank, world)
                local = classify_patches_stain_normalized(slide, coords[lo:hi], net, transform, batch_size, W_est,
                                                          default_target_matrix())
                full = gather_probs(local, len(coords), rank, world)
                probs = full.cpu().numpy() if full is not None else None
            elif world > 1:
                probs = classify_coords_sharded(slide, coords, net, transform, batch_size, rank, world)
            else:
                probs = classify_patches_device(slide, coords, net, transform, batch_size).cpu().numpy()
            if rank == 0:
                write_slide_csv(slide_csv, coords, probs, model_info.config.class_names)
        except Exception as e:  # noqa: BLE001 - slide-granular isolation, as upstream :268-280
            failed_inference.append(stem)
            logger.error(f"Failed inference on {wsi_path}", exc_info=e)
    return failed_patching, failed_inference


class HostPatchClassifier:
    """Host-buffer form of the classification loop (run_inference.py:539-556).

    ``submit(rgba, probs)`` takes a pinned uint8 CPU tensor ``(n, ps, pitch_px, 4)`` holding the
    RGBA pixels ``read_region`` returned for ``n`` patches (rows padded to a multiple of 4 pixels)
    and a pinned fp32 CPU tensor ``(n, K)`` that receives the probabilities after ``sync()``.
    H2D copy, fused preprocess, forward and D2H copy run inside ``libwsi_b200`` on its own two
    streams with ``depth`` batches in flight.
    """

    def __init__(self, net, patch_size: int, transform, max_batch: int, depth: int = 2, host_channels: int = 4):
        """``host_channels=3``: the host batches are dense RGB ``(n, ps, ps, 3)`` -- what
        ``read_region(...).convert("RGB")`` yields (data.py:276-281) -- 25 % fewer bytes over PCIe, expanded to the RGBA
        stack on the device; ``4`` (default): RGBA rows padded to a multiple of 4 pixels."""
        import ctypes as C

        torch = _cabi.require_cuda()
        self.net, self.patch_size, self.max_batch = net, int(patch_size), int(max_batch)
        self.pitch_px = (self.patch_size + 3) // 4 * 4
        self.S = transform.out_size(patch_size)
        self._h = C.c_void_p()
        mean = _cabi.float_array(transform.mean) if transform.mean is not None else None
        std = _cabi.float_array(transform.std) if transform.std is not None else None
        with torch.cuda.device(net.device):
            _cabi.call("wsi_hostpipe_create", net._h, self.patch_size, self.pitch_px * 4, self.S, mean, std,
                       self.max_batch, int(depth), C.byref(self._h))
            self.host_channels = int(host_channels)
            if self.host_channels != 4:
                _cabi.call("wsi_hostpipe_set_host_channels", self._h, self.host_channels)

    def alloc_host_batch(self, n: int = None):
        """Pinned ``(n, ps, pitch_px, 4)`` uint8 and ``(n, K)`` fp32 host tensors."""
        import torch

        n = self.max_batch if n is None else int(n)
        shape = (n, self.patch_size, self.patch_size, 3) if self.host_channels == 3 else \
            (n, self.patch_size, self.pitch_px, 4)
        return (torch.zeros(shape, dtype=torch.uint8).pin_memory(),
                torch.zeros((n, self.net.num_classes), dtype=torch.float32).pin_memory())

    def submit(self, rgba, probs):
        n = int(rgba.shape[0])
        assert rgba.dtype.is_floating_point is False and rgba.is_contiguous() and not rgba.is_cuda
        assert tuple(rgba.shape[1:]) == ((self.patch_size, self.patch_size, 3) if self.host_channels == 3
                                         else (self.patch_size, self.pitch_px, 4))
        assert probs.is_contiguous() and not probs.is_cuda and probs.numel() >= n * self.net.num_classes
        _cabi.call("wsi_hostpipe_submit", self._h, rgba.data_ptr(), n, probs.data_ptr())

    def sync(self):
        _cabi.call("wsi_hostpipe_sync", self._h)

    @property
    def compute_stream(self):
        import torch

        return torch.cuda.ExternalStream(int(_cabi.load().wsi_hostpipe_compute_stream(self._h)), device=self.net.device)

    def bytes_copied(self):
        import ctypes as C

        a, b = C.c_longlong(), C.c_longlong()
        _cabi.call("wsi_hostpipe_bytes", self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def close(self):
        if self._h:
            _cabi.load().wsi_hostpipe_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
