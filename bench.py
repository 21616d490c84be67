#!/usr/bin/env python
"""bench.py -- patches/s of the WSInsight patch-classification hot path on B200.

Workload (BASELINE.json configs[1]): breast-tumor-resnet34.tcga-brca on a synthetic
100 000 x 80 000 slide at 0.5 um/px, bf16, batch 1024.  One "step" = one batch of 1024
in-tissue patches: fused crop/resize/normalise kernel + ResNet34 forward (softmax in the tail
kernel).

  value  patches/s with the slide resident in HBM (device-timed, CUDA events, max over ranks)
  e2e    patches/s through the host-buffer entry (wsi_hostpipe_*): every step copies the
         batch's RGBA pixels from pinned host memory and reads the probabilities back
  roofline      the tcgen05 convolution kernel against the measured bf16 peak
  cpu_baseline  the reference's CPU path (PIL/torchvision transform + fp32 TorchScript ResNet34)
                timed on the same box's host cores on a bounded sample (rank 0, N=1)

``--impl reference`` times only that CPU path (see ``reference_arm``).
Launch: ``python bench.py [--gpus N --steps K --warmup W]`` or, for N>1, under torchrun.
"""

from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

_JSON_FD = None        # original stdout when fd 1 had to be pointed at stderr (see b200_arm)


def emit_json(line: dict) -> None:
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

METRIC = "patches/sec (ResNet34 224^2, bf16)"
UNIT = "patches/s"
RESNET34_GFLOP = 7.326  # convs + FC, 2*MACs, per 224^2 patch


class ModelSpec:
    """The two zoo models of BASELINE.json: configs[1] (default, the metric's model) and configs[2]."""

    def __init__(self, name):
        from oracle import forward as oforward

        self.name = name
        if name == "resnet34":
            self.cfg = oforward.BREAST_RESNET34_CONFIG
            # BN-calibrated random init: probabilities are NOT saturated (the default-init net is)
            self.make = lambda: oforward.make_resnet34_calibrated(2, 0)
            self.gflop, self.metric = RESNET34_GFLOP, METRIC
            self.title, self.weights = ("breast-tumor-resnet34.tcga-brca",
                                        "torchvision resnet34(num_classes=2), BN statistics calibrated, BN folded")
        else:
            from oracle import inceptionv4 as oinc

            self.cfg = oinc.TIL_INCEPTIONV4_CONFIG
            self.make = lambda: oinc.make_inceptionv4(2, 0)
            self.gflop, self.metric = 24.505, "patches/sec (InceptionV4 299^2, bf16)"
            self.title, self.weights = "pancancer-lymphocytes-inceptionv4.tcga", "Cadene-layout InceptionV4 without BatchNorm"
        self.S = int(self.cfg["transform"][0]["arguments"]["size"])

    def patch_px(self, mpp):
        return int(round(self.cfg["patch_size_pixels"] * self.cfg["spacing_um_px"] / mpp))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--model", default="resnet34", choices=["resnet34", "inceptionv4"],
                    help="resnet34 = BASELINE configs[1] (the metric); inceptionv4 = configs[2]")
    ap.add_argument("--slide", default="100000x80000", help="WxH of the synthetic slide (level 0 pixels)")
    ap.add_argument("--mpp", type=float, default=0.5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--per-op", action="store_true", help="print the per-layer event timings to stderr")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--sustained-seconds", type=float, default=3.2,
                    help="length of the second, long timed window the sustained roofline is taken over (N=1)")
    ap.add_argument("--quick", action="store_true",
                    help="only value / e2e / roofline (skip library comparator, cohort, InceptionV4, parity, strong scaling)")
    ap.add_argument("--parity-patches", type=int, default=10240)
    return ap.parse_args()


def workload_name(w, h, mpp, batch, spec=None):
    if spec is None or spec.name == "resnet34":
        return (f"breast-tumor-resnet34.tcga-brca on a synthetic {w}x{h} slide at {mpp} um/px, "
                f"patch 350px@0.25um -> 224^2, batch {batch}")
    return (f"{spec.title} on a synthetic {w}x{h} slide at {mpp} um/px, patch {spec.cfg['patch_size_pixels']}px@"
            f"{spec.cfg['spacing_um_px']}um -> {spec.S}^2, batch {batch}")


# ------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region (B200_PROFILING.md "clocks" line)
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def mark(self):
        return time.perf_counter()

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for t, line in self.lines:
            if not (t0 <= t <= t1 + 0.15):
                continue
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own CPU path for this hot path (test/bench infrastructure: oracle/)
# ------------------------------------------------------------------------------------------------
def host_patches(coords_xy, ps, slide_w, slide_h, seed):
    """Regenerate RGBA patch windows of the synthetic slide on the HOST (same counter-based
    generator as the device slide; zero padded outside the slide like read_region)."""
    import numpy as np

    from wsinsight_b200.wsi import synth_window_host

    out = np.zeros((len(coords_xy), ps, ps, 4), dtype=np.uint8)
    for i, (x, y) in enumerate(coords_xy):
        x, y = int(x), int(y)
        x1, y1 = min(x + ps, slide_w), min(y + ps, slide_h)
        if x1 > x and y1 > y:
            out[i, : y1 - y, : x1 - x] = synth_window_host(x, y, x1 - x, y1 - y, slide_w, slide_h, seed)
    return out


class CpuReference:
    """PIL convert + torchvision Compose(Resize, ToTensor, Normalize) + frozen TorchScript ResNet34
    fp32 + CPU softmax, batch 32 (wsinsight/modellib/data.py:270-314, transforms.py:22-38,
    run_inference.py:288-299,537-560), all host threads."""

    def __init__(self, batch=None, spec=None):
        import torch

        from oracle import forward as oforward

        self.cores = len(os.sched_getaffinity(0))
        torch.set_num_threads(self.cores)
        # reference default batch is 32 (cli/run.py:369-376); many-core hosts get a larger batch so
        # that the intra-op thread pool of the fp32 forward is not starved
        self.batch = batch if batch is not None else (32 if self.cores <= 16 else 128)
        spec = spec or ModelSpec("resnet34")
        self.cfg = spec.cfg
        model = spec.make()
        self.ts = torch.jit.freeze(torch.jit.script(model))
        from concurrent.futures import ThreadPoolExecutor

        self.pool = ThreadPoolExecutor(max_workers=min(self.cores, 8))   # reference default: min(cpus, 8) workers
        tf = self.cfg["transform"]
        self.size = tf[0]["arguments"]["size"]
        self.mean, self.std = tf[2]["arguments"]["mean"], tf[2]["arguments"]["std"]

    def step(self, rgba_batch):
        """One batch: per-patch PIL transform on worker threads + forward + softmax."""
        import torch

        from oracle import forward as oforward
        from oracle import preprocess

        fn = lambda p: preprocess.preprocess_lib(p, self.size, self.mean, self.std)  # noqa: E731
        x = torch.stack(list(self.pool.map(fn, rgba_batch)), 0)   # PIL releases the GIL: ~DataLoader workers
        with torch.no_grad():
            return oforward.head(self.ts(x))


def sample_grid_coords(slide_w, slide_h, ps, n, seed=0):
    """Seeded sample of patch-grid positions (x-major grid of pipeline.py:359-382, zero overlap)."""
    import numpy as np

    half = round(ps / 2)
    ni, nj = len(range(half, slide_w, ps)), len(range(half, slide_h, ps))
    rng = np.random.default_rng(seed)
    idx = rng.choice(ni * nj, size=min(n, ni * nj), replace=False)
    idx.sort()
    return np.stack([(idx // nj) * ps, (idx % nj) * ps], 1).astype(np.int64)


def time_cpu_reference(ref, coords_xy, ps, slide_w, slide_h, seed, steps, warmup, budget_s=None):
    """Time `steps` batches of ref.batch patches after `warmup` (stop early once budget_s is spent)."""
    b = ref.batch
    need = (steps + warmup) * b
    reps = -(-need // len(coords_xy))
    import numpy as np

    coords = np.concatenate([coords_xy] * reps)[:need]
    times = []
    for s in range(steps + warmup):
        patches = host_patches(coords[s * b:(s + 1) * b], ps, slide_w, slide_h, seed)
        t0 = time.perf_counter()
        ref.step(patches)
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
            if budget_s is not None and sum(times) >= budget_s and len(times) >= 2:
                break
    return times


def reference_arm(args, rank, world):
    """`--impl reference`: the reference's CPU implementation of the path on this box's host cores.

    The reference package itself cannot be imported in this image; this executes the
    same PIL / torchvision / TorchScript calls its data path and classification loop make
    (oracle/preprocess.py, oracle/forward.py).  Rank 0 alone runs."""
    if rank != 0:
        return
    w, h = (int(v) for v in args.slide.lower().split("x"))
    spec = ModelSpec(args.model)
    ps = spec.patch_px(args.mpp)
    ref = CpuReference(spec=spec)
    coords = sample_grid_coords(w, h, ps, (args.steps + args.warmup) * ref.batch, seed=0)
    t_all0 = time.perf_counter()
    times = time_cpu_reference(ref, coords, ps, w, h, 0, args.steps, args.warmup)
    total = sum(times)
    val = len(times) * ref.batch / total
    sample = (f"{len(times)} timed batches of {ref.batch} patches ({ps}px RGBA -> {spec.S}^2) at seeded grid positions of "
              f"the {w}x{h} synthetic slide, regenerated on the host; PIL/torchvision transform on min(cores,8) worker threads + frozen "
              f"TorchScript {spec.name} fp32 + softmax, {ref.cores} threads")
    line = {
        "impl": "reference", "metric": spec.metric, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": len(times),
        "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(w, h, args.mpp, args.batch, spec), "batch_per_step": ref.batch,
                   "note": "bounded sample of the same workload; CPU batch 32 (reference default) or 128 on >16 cores"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": ref.cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.perf_counter() - t_all0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def load_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": float(d["hbm_gbs"]), "tf_burst": float(d["bf16_tflops"]),
                "tf_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback"}


class SlideInfo:
    """Geometry-only stand-in for a slide whose thumbnail is already on the device (cohort stage)."""

    def __init__(self, w, h, mpp, name):
        self.width, self.height, self.mpp, self.name = int(w), int(h), float(mpp), name

    @property
    def dimensions(self):
        return (self.width, self.height)


def conv_summary(ops, nfwd, B):
    conv_ms = sum(o["ms_total"] for o in ops if o["kind"] == "conv")
    conv_fl = sum(o["flops_per_image"] for o in ops if o["kind"] == "conv")
    n_conv = sum(1 for o in ops if o["kind"] == "conv")
    other_ms = sum(o["ms_total"] for o in ops if o["kind"] != "conv")
    return dict(conv_ms=conv_ms, conv_fl=conv_fl, n_conv=n_conv, other_ms=other_ms,
                achieved=conv_fl * B * nfwd / max(conv_ms * 1e-3, 1e-12) / 1e12)


def stage_library_baseline(spec, B, dev, steps=8):
    """The SAME frozen TorchScript model through PyTorch/cuDNN on this GPU ('the kernel to beat on the
    same box'): fp32 as the reference runs it (run_inference.py:542-545; torch defaults, i.e. TF32 convolutions allowed),
    strict IEEE fp32, and bf16 channels_last.  CUDA-event timed, pre-staged input and with .to(device) / .cpu() in the loop."""
    import torch

    from oracle import forward as oforward

    out = {}
    S = spec.S
    xh = torch.randn((B, 3, S, S)).pin_memory()

    def run(model, x_dev, pre):
        with torch.no_grad():
            for _ in range(3):
                model(x_dev)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                model(x_dev)
            e1.record()
            torch.cuda.synchronize()
            staged = B * steps / (e0.elapsed_time(e1) * 1e-3)
            t0 = time.perf_counter()
            for _ in range(steps):
                p = oforward.head(model(pre(xh.to(dev, non_blocking=True))).float().cpu())   # reference loop :542-551
            torch.cuda.synchronize()
            e2e = B * steps / (time.perf_counter() - t0)
        return staged, e2e

    tf32_c, tf32_m = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    try:
        m = spec.make().to(dev)
        ts = torch.jit.freeze(torch.jit.script(m))
        xd = xh.to(dev)
        a, b = run(ts, xd, lambda t: t)
        out["fp32_torch_default"] = {"patches_per_s": a, "e2e_patches_per_s": b,
                                     "note": f"cudnn.allow_tf32={tf32_c} (torch default, as the reference runs)"}
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        a, b = run(ts, xd, lambda t: t)
        out["fp32_ieee"] = {"patches_per_s": a, "e2e_patches_per_s": b}
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32_c, tf32_m
        mb = spec.make().to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
        tsb = torch.jit.freeze(torch.jit.script(mb))
        cl = lambda t: t.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)   # noqa: E731
        a, b = run(tsb, cl(xd), cl)
        out["bf16_channels_last"] = {"patches_per_s": a, "e2e_patches_per_s": b}
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = tf32_c, tf32_m
    out["batch"] = B
    out["h2d_bytes_per_step_e2e"] = B * 3 * S * S * 4
    out["what"] = ("frozen TorchScript " + spec.name + " through PyTorch/cuDNN on the same GPU, CUDA-event timed over "
                   f"{steps} batches after 3 warm-up batches; e2e = pinned fp32 host batch .to(device) + forward + "
                   ".cpu() + CPU softmax per step (the reference's loop)")
    del xh
    torch.cuda.empty_cache()
    return out


def stage_cohort64(dev, peaks, cfg, n_slides=64, check=4):
    """BASELINE config 4: tissue segmentation + patchify of a 64-slide cohort in ONE batched launch sequence.

    Slides: seeds 0..63, the four SURVEY sizes (thumbnails 2048x2048, 2048x1638, 1536x2048, 2048x1024), each generated at
    level 0 (16x the thumbnail), thumbnailed with the Pillow-exact kernel and freed.  Masks of `check` slides and their
    coordinates are compared with the CPU oracle."""
    import numpy as np
    import torch

    from oracle import patchgrid, segment
    from wsinsight_b200.patchlib import segment_and_patch_slides_device
    from wsinsight_b200.wsi import DeviceSlide

    sizes = [(32768, 32768), (32768, 26208), (24576, 32768), (32768, 16384)]
    infos, thumbs = [], []
    t0 = time.perf_counter()
    for sd in range(n_slides):
        w, h = sizes[sd % 4]
        sl = DeviceSlide.synthetic(w, h, mpp=0.5, seed=sd, device=dev)
        thumbs.append(sl.thumbnail_device((2048, 2048)).clone())
        infos.append(SlideInfo(w, h, 0.5, f"cohort_{sd:02d}"))
        del sl
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t0
    ms = []
    res = None
    for _ in range(4):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = segment_and_patch_slides_device(infos, cfg["patch_size_pixels"], cfg["spacing_um_px"], thumbs=thumbs)
        torch.cuda.synchronize()
        ms.append((time.perf_counter() - t0) * 1e3)
    npx = sum(int(t.shape[0]) * int(t.shape[1]) for t in thumbs)
    npatch = sum(0 if r.coords is None else int(r.coords.shape[0]) for r in res)
    alg = 4 * npx + 8 * npatch                   # 3*Ht*Wt read + Ht*Wt written + 8 B per kept patch
    best = min(ms[1:])
    ok_mask = ok_coords = 0
    for i in list(range(0, n_slides, max(1, n_slides // check)))[:check]:
        r = res[i]
        want = segment.segment_tissue(thumbs[i].cpu().numpy(), 7, 7, 6, r.min_object_size_px, r.min_hole_size_px)
        ok_mask += bool(np.array_equal(r.mask.cpu().numpy().astype(bool), want))
        cent, _, _ = patchgrid.grid_centroids(infos[i].width, infos[i].height, r.patch_size, r.half_patch_size)
        keep = patchgrid.raster_rule(segment.outermost_restated(want), cent, infos[i].width, infos[i].height)
        got = r.coords.cpu().numpy().astype(np.int64) if r.coords is not None else np.zeros((0, 2), np.int64)
        ok_coords += bool(np.array_equal(got, cent[keep] - r.half_patch_size))
    return {"slides": n_slides, "ms_total": best, "ms_per_slide": best / n_slides, "ms_runs": ms,
            "thumbnail_pixels": npx, "in_tissue_patches": npatch, "algorithmic_bytes": alg,
            "achieved_gbs": alg / (best * 1e-3) / 1e9, "peak_gbs": peaks["hbm_gbs"],
            "frac": alg / (best * 1e-3) / 1e9 / peaks["hbm_gbs"], "bound": "hbm (latency-bound at this size)",
            "timing": "host wall clock around the whole cohort call incl. the one device->host read of counts and flags",
            "oracle_checked_slides": check, "masks_equal_oracle": ok_mask, "coords_equal_oracle": ok_coords,
            "slide_generation_plus_thumbnail_s": t_gen}


def stage_inceptionv4(slide, dev, peaks, B, seconds, rank=0, world=1):
    """BASELINE config 3: pancancer-lymphocytes-inceptionv4.tcga, 100 um patches of the same slide sharded over the ranks
    (contiguous x-bands of the patch list), >= `seconds` of back-to-back steps per rank, MAX over ranks."""
    import torch
    import torch.distributed as dist

    from wsinsight_b200.modellib.data import PreprocessPlan
    from wsinsight_b200.modellib.engine import B200Network
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from wsinsight_b200.patchlib import segment_and_patch_slides_device

    from wsinsight_b200.modellib.run_inference import shard_range

    spec = ModelSpec("inceptionv4")
    cfg = spec.cfg
    setup_err = None
    try:
        # everything that can fail on ONE rank happens before the first collective of this stage
        res = segment_and_patch_slides_device([slide], cfg["patch_size_pixels"], cfg["spacing_um_px"])[0]
        n_all = int(res.coords.shape[0])
        lo, hi = shard_range(n_all, rank, world)
        coords = res.coords[lo:hi].contiguous()
        ps = res.patch_size
        nb = int(coords.shape[0]) // B
        assert nb >= 1, "not enough patches for one batch"
        net = B200Network.from_torchscript(torch.jit.freeze(torch.jit.script(spec.make())), 2, input_size=spec.S,
                                           max_batch=B, device=dev)
        transform = make_compose_from_transform_config(cfg["transform"])
        lay = net.input_layout
        plan = PreprocessPlan(slide, ps, transform, lay["out_H"], lay["out_W"], lay["pad_top"], lay["pad_left"],
                              lay["plane_stride_px"], lay["channels"])
        xin = net.alloc_input(B)
        probs = torch.empty((B, 2), dtype=torch.float32, device=dev)

        def step(i):
            b = i % nb
            plan.run(coords[b * B:(b + 1) * B], xin)
            net.forward_into(xin, probs)

        for i in range(3):
            step(i)
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        setup_err = e
    if world > 1:
        bad = torch.tensor([1 if setup_err is not None else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(bad, op=dist.ReduceOp.MAX)
        if int(bad.item()):
            raise RuntimeError(f"InceptionV4 stage skipped on all ranks: setup failed on a rank ({setup_err})")
    elif setup_err is not None:
        raise setup_err
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(4):
        step(3 + i)
    e1.record()
    torch.cuda.synchronize()
    per = e0.elapsed_time(e1) / 4
    n = max(8, int(seconds * 1e3 / per) + 1)
    if world > 1:                      # every rank times the same number of steps
        nn = torch.tensor([n], dtype=torch.int64, device=dev)
        dist.all_reduce(nn, op=dist.ReduceOp.MAX)
        n = int(nn.item())
        torch.cuda.synchronize()
        dist.barrier()
    e0.record()
    for i in range(n):
        step(7 + i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        tm = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        ms = float(tm.item())
    value = world * n * B / (ms * 1e-3)
    # conv time and traffic figures from a separate short pass with per-op events
    net.set_profiling(8)
    for i in range(8):
        step(i)
    ops, nfwd = net.profile()
    net.set_profiling(0)
    cs = conv_summary(ops, nfwd, B)
    # algorithmic HBM bytes per image of the conv launches: every conv reads its input once and writes its output once
    alg = float(_cabi_call_double("wsi_net_conv_bytes_per_image", net)) if hasattr(net, "_h") else None
    out = {"metric": spec.metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": n, "seconds": ms * 1e-3, "batch": B,
           "ms_per_step": ms / n, "patch_px": ps, "in_tissue_patches": n_all,
           "conv_tflops": cs["achieved"], "conv_frac_of_sustained": cs["achieved"] / peaks["tf_sustained"],
           "conv_share_of_step": cs["conv_ms"] / nfwd / (ms / n), "pool_head_ms_per_step": cs["other_ms"] / nfwd,
           "tensor_pipe_frac_of_step": value / world * spec.gflop * 1e9 / (peaks["tf_sustained"] * 1e12),
           "conv_launches": cs["n_conv"], "launches_per_step": net.num_launches + 1,
           "conv_algorithmic_bytes_per_launch_avg": alg * B / cs["n_conv"] if alg else None,
           "conv_dram_bytes_per_launch_avg_ncu": ncu_traffic("inceptionv4"),
           "workload": workload_name(slide.width, slide.height, slide.mpp, B, spec)}
    plan.close()
    net.close()
    torch.cuda.empty_cache()
    return out


def ncu_traffic(model):
    """DRAM bytes per conv launch (read + write) from the committed ncu capture of this command, newest round first."""
    for name in ("r02_ncu_conv_traffic.json", "r01_ncu_conv_traffic.json"):
        try:
            v = json.load(open(ROOT / "profiles" / name)).get(model, {}).get("dram_bytes_per_launch_avg")
            if v:
                return v
        except Exception:  # noqa: BLE001
            continue
    return None


def _cabi_call_double(name, net):
    from wsinsight_b200 import _cabi

    fn = getattr(_cabi.load(), name, None)
    return fn(net._h) if fn is not None else 0.0


def stage_strong_scaling(slide, thumb, spec, net, transform, B, rank, world, dev):
    """ONE slide classified by all ranks (the product's multi-GPU path): seg + patchify on every rank (it is ~1 ms and
    keeps the coordinates implicit), contiguous x-band shard per rank, gather of the (N_r, K) shards over NCCL, rank 0
    holds the (N, K) host array.  Wall clock around everything, max over ranks."""
    import torch
    import torch.distributed as dist

    from wsinsight_b200.modellib.run_inference import classify_slide_sharded

    cfg = spec.cfg
    n_total = 0
    keep = {}

    def timed(result_path, reps):
        nonlocal n_total
        runs = []
        for it in range(reps):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            coords_np, probs_np = classify_slide_sharded(slide, cfg["patch_size_pixels"], cfg["spacing_um_px"], net,
                                                         transform, B, rank=rank, world=world, thumb=thumb,
                                                         result_path=result_path)
            torch.cuda.synchronize()
            dt = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            runs.append(float(dt.item()))
            if rank == 0:
                n_total = len(coords_np)
                assert probs_np.shape == (n_total, net.num_classes)
                keep[result_path] = probs_np
        return runs

    runs = timed("nccl", 2)
    peer = None
    if world > 1 and os.environ.get("WSI_B200_NO_PEER_STAGE") != "1":
        # the same slide with the tail kernel storing into a rank-0 buffer mapped through CUDA IPC (no collective)
        try:
            pr = timed("peer", 2)
            peer = {"ms_per_slide": min(pr), "ms_runs": pr,
                    "identical_to_nccl_gather": bool(rank != 0 or (keep["peer"] == keep["nccl"]).all())}
        except Exception as e:  # noqa: BLE001
            peer = {"error": f"{type(e).__name__}: {e}"}
    return {"ms_per_slide": min(runs), "ms_runs": runs, "patches": n_total, "ranks": world, "peer_store_result_path": peer,
            "patches_per_s": n_total / (min(runs) * 1e-3) if n_total else None,
            "includes": "seg + patchify (every rank), fused preprocess + forward of the rank's x-band, "
                        + ("NCCL gather of (N_r, K) shards to rank 0, " if world > 1 else "") + "(N, K) host array",
            "slide_replica_per_rank": True}


def b200_arm(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge

    ge.build()
    from wsinsight_b200 import _cabi
    from wsinsight_b200.modellib.data import PreprocessPlan
    from wsinsight_b200.modellib.engine import B200Network
    from wsinsight_b200.modellib.run_inference import HostPatchClassifier, gather_probs, shard_range
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from wsinsight_b200.patchlib import segment_and_patch_slides_device
    from wsinsight_b200.wsi import DeviceSlide

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        # NCCL writes its version banner to file descriptor 1 from C code: point fd 1 at stderr for the rest of the
        # run and keep the original stdout for the one JSON line of the contract
        global _JSON_FD
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    lib = _cabi.load()
    peaks = load_peaks()
    W, H = (int(v) for v in args.slide.lower().split("x"))
    B, K, Wm = args.batch, args.steps, args.warmup
    spec = ModelSpec(args.model)
    cfg = spec.cfg
    transform = make_compose_from_transform_config(cfg["transform"])
    extras = world == 1 and not args.quick

    def ev():
        return torch.cuda.Event(enable_timing=True)

    # ---- setup: slide in HBM, tissue segmentation + patchify (timed as its own stage) ----
    t0 = time.perf_counter()
    slide = DeviceSlide.synthetic(W, H, mpp=args.mpp, seed=0, device=dev)
    torch.cuda.synchronize()
    t_gen = time.perf_counter() - t0
    # thumbnail: Pillow-exact LANCZOS chain, the only stage that reads every level-0 pixel
    thumb_ms = []
    for i in range(3):
        slide._thumb_cache.clear()
        e0, e1 = ev(), ev()
        e0.record()
        thumb = slide.thumbnail_device((2048, 2048))
        e1.record()
        torch.cuda.synchronize()
        thumb_ms.append(e0.elapsed_time(e1))
    seg_ms, seg_wall = [], []
    for i in range(5):
        e0, e1 = ev(), ev()
        torch.cuda.synchronize()
        tw = time.perf_counter()
        e0.record()
        res = segment_and_patch_slides_device([slide], cfg["patch_size_pixels"], cfg["spacing_um_px"], thumbs=[thumb])[0]
        e1.record()
        torch.cuda.synchronize()
        seg_wall.append((time.perf_counter() - tw) * 1e3)
        seg_ms.append(e0.elapsed_time(e1))
    ps = res.patch_size
    coords_all = res.coords
    n_total = int(coords_all.shape[0])
    lo, hi = shard_range(n_total, rank, world)
    coords = coords_all[lo:hi].contiguous()
    n_local = int(coords.shape[0])
    assert n_local >= B, f"rank {rank}: only {n_local} in-tissue patches for batch {B}"
    nb_local = n_local // B

    model = spec.make()
    net = B200Network.from_torchscript(torch.jit.freeze(torch.jit.script(model)), 2, input_size=spec.S, max_batch=B,
                                       device=dev)
    lay = net.input_layout
    plan = PreprocessPlan(slide, ps, transform, lay["out_H"], lay["out_W"], lay["pad_top"], lay["pad_left"],
                          lay["plane_stride_px"], lay["channels"])
    xin = net.alloc_input(B)
    probs = torch.empty((nb_local * B, 2), dtype=torch.float32, device=dev)

    def step(i):
        b = i % nb_local
        plan.run(coords[b * B:(b + 1) * B], xin)
        net.forward_into(xin, probs[b * B:(b + 1) * B])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- pass A: the contract's timed region (device-resident, NO per-op events inside) ----
    for i in range(Wm):
        step(i)
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    barrier()
    launches0 = int(lib.wsi_launch_count())
    c0 = sampler.mark()
    e0, e1 = ev(), ev()
    e0.record()
    for i in range(K):
        step(Wm + i)
    e1.record()
    barrier()
    c1 = sampler.mark()
    launches = int(lib.wsi_launch_count()) - launches0
    ms = e0.elapsed_time(e1)
    tms = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
    ms_max = float(tms.item())
    value = world * K * B / (ms_max * 1e-3)

    # ---- e2e: host buffers through wsi_hostpipe_*; the per-rank results are gathered on rank 0 inside the region ----
    # (timed right after pass A, before the long sustained pass drives the GPU into its power cap: value and e2e see
    # the same thermal state)
    e2e = None
    e2e_rgba = None

    def run_e2e(host_channels):
        """K steps through wsi_hostpipe_* from pinned host patches (3: dense RGB as read_region().convert("RGB") yields,
        4: padded RGBA rows), probabilities back on the host, per-rank results gathered on rank 0 inside the region."""
        depth = 3
        hp = HostPatchClassifier(net, ps, transform, B, depth=depth, host_channels=host_channels)
        nbuf = min(4, nb_local)
        host = []
        for j in range(nbuf):
            px, hprobs = hp.alloc_host_batch(B)
            cj = coords[j * B:(j + 1) * B].cpu().numpy()
            for i, (x, y) in enumerate(cj):
                x1, y1 = min(x + ps, W), min(y + ps, H)
                px[i, : y1 - y, : x1 - x].copy_(slide.rgba[y:y1, x:x1, :host_channels])
            host.append((px, hprobs))
        torch.cuda.synchronize()

        def gather_results():
            # what a multi-rank caller does with its host results: one (N_r, K) shard per rank -> rank 0 (NCCL)
            if world == 1:
                return None
            shard = torch.cat([h[1] for h in host], 0).to(dev, non_blocking=True)
            bufs = [torch.empty_like(shard) for _ in range(world)] if rank == 0 else None
            dist.gather(shard, bufs, dst=0)
            return torch.cat(bufs, 0).cpu() if rank == 0 else None

        for i in range(max(Wm, 3)):
            hp.submit(*host[i % nbuf])
        hp.sync()
        gather_results()
        h2d0, d2h0 = hp.bytes_copied()
        barrier()
        cs_ = hp.compute_stream
        s0, s1 = ev(), ev()
        ec0 = sampler.mark()
        tw0 = time.perf_counter()
        s0.record(cs_)
        for i in range(K):
            hp.submit(*host[(Wm + i) % nbuf])
        s1.record(cs_)
        hp.sync()
        gather_results()
        tw1 = time.perf_counter()
        barrier()
        ec1 = sampler.mark()
        h2d1, d2h1 = hp.bytes_copied()
        # the timed region ends when the last probabilities are on the host (rank 0's host for N > 1): take the
        # larger of the device-event span and the host wall clock around submit..sync..gather
        e_ms = max(s0.elapsed_time(s1), (tw1 - tw0) * 1e3)
        t2 = torch.tensor([e_ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e_ms = float(t2.item())
        ref_p = probs[: nbuf * B].cpu() if Wm + K >= nbuf else None
        out = {"value": world * K * B / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": (h2d1 - h2d0) // K,
               "d2h_bytes_per_step": (d2h1 - d2h0) // K, "ms_per_step": e_ms / K, "in_flight": depth,
               "host_buffers": ("dense RGB u8 (n, ps, ps, 3): what read_region(...).convert('RGB') yields" if host_channels == 3
                                else "RGBA u8 rows padded to 4 pixels: what read_region returns"),
               "collective_in_region": "one NCCL gather of the per-rank (N_r, K) results to rank 0" if world > 1 else "none (1 rank)"}
        out["clocks"] = sampler.summary(ec0, ec1)
        if ref_p is not None:
            got = torch.cat([h[1] for h in host], 0)
            out["max_abs_diff_vs_resident_path"] = float((got - ref_p).abs().max().item())
        hp.close()
        return out

    if not args.no_e2e:
        e2e = run_e2e(3)
        e2e_rgba = run_e2e(4)

    # ---- pass B: the same K steps again with per-op CUDA events (roofline of the conv kernels, short window) ----
    net.set_profiling(K)
    pp_ev = [(ev(), ev()) for _ in range(K)]
    barrier()
    b0 = sampler.mark()
    for i in range(K):
        b = (Wm + i) % nb_local
        pp_ev[i][0].record()
        plan.run(coords[b * B:(b + 1) * B], xin)
        pp_ev[i][1].record()
        net.forward_into(xin, probs[b * B:(b + 1) * B])
    barrier()
    b1 = sampler.mark()
    ops, nfwd = net.profile()
    net.set_profiling(0)
    pp_ms = sum(a.elapsed_time(b) for a, b in pp_ev)
    cs = conv_summary(ops, nfwd, B)

    # ---- pass C (N = 1): >= 3 s of back-to-back steps: sustained clocks, sustained roofline ----
    sustained = None
    if world == 1 and args.sustained_seconds > 0:
        n_sus = max(K, int(args.sustained_seconds * 1e3 / (ms / K)) + 1)
        # C1: throughput without events
        barrier()
        s0m = sampler.mark()
        e0, e1 = ev(), ev()
        e0.record()
        for i in range(n_sus):
            step(i)
        e1.record()
        torch.cuda.synchronize()
        s1m = sampler.mark()
        sus_ms = e0.elapsed_time(e1)
        # C2: conv time over an equally long window with per-op events (the clocks stay in the sustained regime)
        net.set_profiling(n_sus)
        q0m = sampler.mark()
        for i in range(n_sus):
            step(i)
        torch.cuda.synchronize()
        q1m = sampler.mark()
        ops_s, nfwd_s = net.profile()
        net.set_profiling(0)
        css = conv_summary(ops_s, nfwd_s, B)
        v_sus = n_sus * B / (sus_ms * 1e-3)
        sustained = {"seconds": sus_ms * 1e-3, "steps": n_sus, "value": v_sus, "unit": UNIT,
                     "achieved": css["achieved"], "peak": peaks["tf_sustained"], "unit_roofline": "TFLOP/s",
                     "frac": css["achieved"] / peaks["tf_sustained"],
                     "peak_source": f"{peaks['source']} bf16 sustained (4 s back-to-back matmul)",
                     "tensor_pipe_frac_of_step": v_sus * spec.gflop * 1e9 / (peaks["tf_sustained"] * 1e12),
                     "conv_share_of_step": css["conv_ms"] / nfwd_s / (sus_ms / n_sus),
                     "clocks": sampler.summary(s0m, s1m), "clocks_profiled_window": sampler.summary(q0m, q1m)}

    # ---- density grid of the classified patches: per-rank binning + ONE all-reduce ----
    from wsinsight_b200.insightlib import allreduce_grids, density_grid_device
    nd = min(nb_local, Wm + K) * B
    cdone = coords[:nd]
    xywh = torch.cat([cdone, torch.full_like(cdone, ps)], 1).contiguous()
    pdone = probs[:nd].contiguous()
    dgrid, ddrop = density_grid_device(xywh, pdone, W, H, 2000)      # warm-up (first-call allocations, lazy init)
    allreduce_grids(dgrid)
    d0, d1 = ev(), ev()
    barrier()
    d0.record()
    dgrid, ddrop = density_grid_device(xywh, pdone, W, H, 2000)
    allreduce_grids(dgrid)
    d1.record()
    torch.cuda.synchronize()
    tot = torch.tensor([nd], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    density = {"bin_px": 2000, "grid": list(dgrid.shape), "ms_bin_plus_allreduce": d0.elapsed_time(d1),
               "objects": int(tot.item()), "grid_sum": int(dgrid.sum().item()) + 0,
               "collective": "nccl all_reduce(sum) over ranks" if world > 1 else "none (1 rank)"}

    # ---- one whole slide on all ranks (strong scaling of the product's multi-GPU path) ----
    strong = None
    if not args.quick:
        strong = stage_strong_scaling(slide, thumb, spec, net, transform, B, rank, world, dev)
    iv4 = None
    if not args.quick and args.model == "resnet34":
        try:
            iv4 = stage_inceptionv4(slide, dev, peaks, B, 3.0, rank, world)
        except Exception as e:  # noqa: BLE001 - an auxiliary stage must not lose the headline line
            import traceback

            traceback.print_exc(file=sys.stderr)
            iv4 = {"error": f"{type(e).__name__}: {e}"}
    sampler.stop()
    clocks = sampler.summary(c0, c1)
    clocks_b = sampler.summary(b0, b1)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    if args.per_op:
        for i, o in enumerate(ops):
            tf = o["flops_per_image"] * B * nfwd / max(o["ms_total"], 1e-9) / 1e9
            sys.stderr.write(f"op {i:3d} {o['kind']:8s} ms/step {o['ms_total'] / nfwd:8.4f}  GFLOP/img "
                             f"{o['flops_per_image'] / 1e9:7.4f}  TFLOP/s {tf:8.1f}\n")
    # ---- roofline of the dominant kernels: the tcgen05 convolutions (tensor pipe) ----
    # DRAM bytes per conv launch from the committed `ncu --set full` capture of this same command
    traffic = ncu_traffic(args.model)
    step_ms_b = sum(o["ms_total"] for o in ops) / nfwd + pp_ms / K
    roofline = {"kernel": ("tcgen05 conv kernels: convs::k_conv_stem + convf::k_conv_flat + convk::k_conv_igemm "
                           f"({cs['n_conv']} launches per step)"),
                "bound": "tensor", "achieved": cs["achieved"], "peak": peaks["tf_burst"], "unit": "TFLOP/s",
                "frac": cs["achieved"] / peaks["tf_burst"], "traffic": traffic,
                "algorithmic_bytes_per_launch_avg": float(_cabi_call_double("wsi_net_conv_bytes_per_image", net)) * B / cs["n_conv"],
                "peak_source": (f"{peaks['source']} bf16 BURST (best of 10 cuBLAS 8192^3): this window is "
                                f"{K} steps = {ms_max * 1e-3:.2f} s, shorter than the power-capped regime"),
                "frac_of_sustained_peak": cs["achieved"] / peaks["tf_sustained"],
                "flops_per_launch_avg": cs["conv_fl"] * B / cs["n_conv"],
                "avg_launch_ms": cs["conv_ms"] / (cs["n_conv"] * nfwd),
                "share_of_step": cs["conv_ms"] / nfwd / step_ms_b,
                "timed_in": "a separate pass of the same K steps with per-op CUDA events (pass B), not the value pass",
                "clocks": clocks_b}
    if sustained is not None:
        roofline["sustained"] = sustained
    pp_bytes = ps * ps * 4 + spec.S * spec.S * 3 * 2     # read ps^2 RGBA + write S^2 x 3 bf16
    stages = {
        "preprocess": {"ms_per_step": pp_ms / K, "bound": "hbm", "achieved_gbs": pp_bytes * B * K / (pp_ms * 1e-3) / 1e9,
                       "peak_gbs": peaks["hbm_gbs"], "frac": pp_bytes * B * K / (pp_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                       "bytes_per_patch": pp_bytes},
        "conv": {"ms_per_step": cs["conv_ms"] / nfwd},
        "pool_head": {"ms_per_step": cs["other_ms"] / nfwd},
        "thumbnail": {"ms_per_slide": min(thumb_ms), "bound": "hbm", "bytes": W * H * 4,
                      "achieved_gbs": W * H * 4 / (min(thumb_ms) * 1e-3) / 1e9, "peak_gbs": peaks["hbm_gbs"],
                      "frac": W * H * 4 / (min(thumb_ms) * 1e-3) / 1e9 / peaks["hbm_gbs"],
                      "note": "Pillow-exact reduce + Lanczos-3 (csrc/thumb.cu); includes host tap tables + sync"},
        "density_grid": density,
        "seg_patchify_ms_per_slide": min(seg_ms), "seg_patchify_ms_runs": seg_ms,
        "seg_patchify_wall_ms_per_slide": min(seg_wall),
        "seg_patchify_algorithmic_bytes": 4 * int(thumb.shape[0]) * int(thumb.shape[1]) + 8 * n_total,
        "slide_generation_s": t_gen, "in_tissue_patches": n_total, "patch_size": ps,
    }
    if strong is not None:
        stages["slide_strong_scaling"] = strong
    if iv4 is not None:
        stages["inceptionv4"] = iv4

    line = {
        "metric": spec.metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
        "ms_per_step": ms_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_name(W, H, args.mpp, B, spec), "batch_per_gpu": B, "global_batch": B * world,
                   "patches_sharded": "contiguous x-bands of the x-major patch list, no data-path collective",
                   "l2": "inputs larger than L2 (125 MB of source pixels + GBs of activations per step; "
                         "consecutive steps read different patches)",
                   "weights": f"random-init {spec.weights}, jit.script+freeze"},
        "roofline": roofline, "stages": stages, "clocks": clocks, "gpu_launches": launches,
        "tensor_pipe_frac_of_step": value / world * spec.gflop * 1e9 / (peaks["tf_sustained"] * 1e12),
    }
    if e2e is not None:
        line["e2e"] = e2e
        line["e2e_rgba_host_buffers"] = e2e_rgba

    def guarded(name, fn):
        try:
            return fn()
        except Exception as e:  # noqa: BLE001 - an auxiliary stage must not lose the headline line
            import traceback

            traceback.print_exc(file=sys.stderr)
            return {"error": f"{type(e).__name__}: {e}"}

    if extras:
        if args.model == "resnet34":
            line["gpu_library_baseline"] = guarded("lib", lambda: stage_library_baseline(spec, B, dev))
            lb = line["gpu_library_baseline"]
            if "bf16_channels_last" in lb:
                lb["b200_over_cudnn_bf16"] = value / lb["bf16_channels_last"]["patches_per_s"]
                lb["b200_over_cudnn_fp32_default"] = value / lb["fp32_torch_default"]["patches_per_s"]
                if e2e is not None:
                    lb["b200_e2e_over_cudnn_bf16_e2e"] = e2e["value"] / lb["bf16_channels_last"]["e2e_patches_per_s"]
                    lb["b200_e2e_over_cudnn_fp32_default_e2e"] = e2e["value"] / lb["fp32_torch_default"]["e2e_patches_per_s"]
        stages["cohort64"] = guarded("cohort", lambda: stage_cohort64(dev, peaks, cfg))
        if args.model == "resnet34":

            def parity():
                sys.path.insert(0, str(ROOT / "tests"))
                from parity_harness import run_parity

                return run_parity("resnet34", 2, n_patches=args.parity_patches, batch=B, oracle_mode="cudnn_fp32",
                                  anchor=512, device=local_rank)

            line["parity"] = guarded("parity", parity)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(args, coords_all, ps, W, H, thumb, res, spec)
        if extras:
            line["cpu_baseline"]["config1"] = guarded("c1", lambda: cpu_baseline_config1(args, spec))
    emit_json(line)
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline(args, coords_all, ps, W, H, thumb, res, spec):
    """Bounded CPU sample of the same workload on this box's host cores (rank 0, N=1)."""
    import numpy as np

    from oracle import patchgrid, segment

    ref = CpuReference(spec=spec)
    n = int(coords_all.shape[0])
    rng = np.random.default_rng(0)
    idx = np.sort(rng.choice(n, size=min(n, ref.batch * 40), replace=False))
    coords = coords_all[idx].cpu().numpy().astype(np.int64)
    times = time_cpu_reference(ref, coords, ps, W, H, 0, steps=38, warmup=2, budget_s=args.cpu_seconds)
    val = len(times) * ref.batch / sum(times)
    out = {"value": val, "unit": UNIT, "cores": ref.cores, "kind": "port",
           "sample": (f"{len(times)} timed batches of {ref.batch} in-tissue patches ({ps}px RGBA -> {spec.S}^2, seeded sample of the "
                      f"slide's patch list, pixels regenerated on the host); PIL/torchvision transform on "
                      f"min(cores,8) worker threads + frozen TorchScript {spec.name} fp32 + softmax, {ref.cores} threads; "
                      "transform and forward run back to back rather than overlapped as DataLoader workers would "
                      "(<= 10 % pessimistic for the CPU)"),
           "ms_per_batch": 1e3 * sum(times) / len(times), "batch": ref.batch}
    # seg + patchify on the CPU (cv2 + scipy restatement + contour/point-in-polygon), same thumbnail
    th = thumb.cpu().numpy()
    t0 = time.perf_counter()
    mask = segment.segment_tissue(th, 7, 7, 6, res.min_object_size_px, res.min_hole_size_px)
    t1 = time.perf_counter()
    xy = patchgrid.patch_coordinates(mask, W, H, ps, res.half_patch_size)
    t2 = time.perf_counter()
    out["seg_patchify_ms_per_slide"] = (t2 - t0) * 1e3
    out["seg_ms"] = (t1 - t0) * 1e3
    out["patchify_ms"] = (t2 - t1) * 1e3
    out["seg_patchify_coords_equal_gpu"] = bool(np.array_equal(np.asarray(xy).astype(np.int64),
                                                               coords_all.cpu().numpy().astype(np.int64)))
    return out


def cpu_baseline_config1(args, spec):
    """BASELINE configs[0]: the reference's own CPU-runnable case -- an 8192 x 8192 slide at 0.25 um/px (350-px patches),
    batch 32, CPU path; bounded sample of its 23 x 23 patch grid."""
    ref = CpuReference(batch=32, spec=spec)
    w = h = 8192
    ps = spec.patch_px(0.25)
    coords = sample_grid_coords(w, h, ps, 32 * 12, seed=1)
    times = time_cpu_reference(ref, coords, ps, w, h, 1, steps=10, warmup=2, budget_s=args.cpu_seconds)
    return {"value": len(times) * 32 / sum(times), "unit": UNIT, "cores": ref.cores, "kind": "port", "batch": 32,
            "ms_per_batch": 1e3 * sum(times) / len(times),
            "workload": f"breast-tumor-resnet34.tcga-brca on a synthetic {w}x{h} slide at 0.25 um/px, patch {ps}px -> {spec.S}^2, batch 32",
            "sample": f"{len(times)} timed batches of 32 patches at seeded grid positions (pixels regenerated on the host)"}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        reference_arm(args, rank, world)
        return
    if world == 1 and args.gpus > 1:
        # plain `python bench.py --gpus N`: re-launch under torchrun, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"),
               str(Path(__file__).resolve())] + sys.argv[1:]
        sys.exit(subprocess.call(cmd))
    b200_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
