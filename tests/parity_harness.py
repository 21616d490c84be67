"""End-to-end forward parity of the PRODUCT path on nets whose probabilities are spread (TEST INFRASTRUCTURE).

north_star: "validated against the reference fp32 path within |dp| <= 2e-3 with >= 99.9 % argmax agreement".
A random-init net with saturated outputs passes that vacuously, so this harness

* takes >= 10^4 in-tissue patches of a synthetic slide (segmentation + patchify on the device),
* builds a BN-calibrated ResNet34 (K = 2 softmax / K = 1 sigmoid) or the TIL InceptionV4 with the classifier
  scaled and its bias centred on the workload, so the fp32 probabilities straddle 0.5,
* asserts the spread FIRST (fraction of fp32 p in [0.05, 0.95]),
* pushes the patches through both product entries -- ``classify_patches_device`` (slide resident in HBM) and
  ``wsi_hostpipe_*`` (host RGBA stacks) --
* compares with the oracle: PIL/torchvision preprocess + fp32 torch forward + CPU softmax/sigmoid
  (reference wsinsight/modellib/run_inference.py:537-560, data.py:270-314, transforms.py:22-38),
* and measures the same two numbers for the library path (the same model through PyTorch/cuDNN in bf16
  channels_last fed by the oracle's fp32 tensors) and for the fp32 model with ONLY its conv weights rounded to
  bf16 (the floor no kernel with bf16 MMA operands can beat).

``oracle_mode="cpu"``: every patch through the CPU fp32 forward.  ``"cudnn_fp32"``: the bulk through cuDNN fp32
with TF32 disabled, anchored by the CPU forward on the first ``anchor`` patches (their agreement is reported).
"""

from __future__ import annotations

import copy
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import torch


def _p_first(p):
    """probability of class 0 (K > 1) or the sigmoid output (K == 1) as a 1-D tensor"""
    return p if p.dim() == 1 else p[:, 0]


def _decision(p):
    return (p > 0.5) if p.dim() == 1 else p.argmax(1)


def diff_stats(p, want, band=0.05):
    d = (p.double() - want.double()).abs()
    d1 = d if d.dim() == 1 else d.max(1).values
    agree = (_decision(p) == _decision(want)).double()
    far = (_p_first(want).double() - 0.5).abs() > band
    return {
        "max_abs_dp": float(d1.max()), "p999_abs_dp": float(torch.quantile(d1, 0.999)),
        "p99_abs_dp": float(torch.quantile(d1, 0.99)), "median_abs_dp": float(d1.median()),
        "mean_abs_dp": float(d1.mean()), "frac_within_2e-3": float((d1 <= 2e-3).double().mean()),
        "argmax_agreement": float(agree.mean()),
        f"argmax_agreement_margin_gt_{band}": float(agree[far].mean()) if bool(far.any()) else None,
        "n_margin": int(far.sum()),
    }


def make_model(arch: str, K: int, seed: int = 0):
    from oracle import forward as oforward

    if arch == "resnet34":
        return oforward.make_resnet34_calibrated(K, seed), oforward.BREAST_RESNET34_CONFIG
    from oracle import inceptionv4 as oinc

    return oinc.make_inceptionv4(K, seed, fc_gain=1.0), oinc.TIL_INCEPTIONV4_CONFIG


def center_classifier(model, x_f32, target_logit_std=1.5):
    """Scale the classifier so the decision logit has std ``target_logit_std`` over ``x_f32`` and shift its bias so
    the median patch sits on the decision boundary (p = 0.5): the hardest case for argmax agreement."""
    fc = model.fc if hasattr(model, "fc") else model.last_linear
    with torch.no_grad():
        z = model(x_f32)
        dz = z[:, 0] - z[:, 1] if z.shape[1] > 1 else z[:, 0]
        g = float(target_logit_std / max(float(dz.std()), 1e-9))
        fc.weight.mul_(g)
        fc.bias.mul_(g)
        z = model(x_f32)
        dz = z[:, 0] - z[:, 1] if z.shape[1] > 1 else z[:, 0]
        m = float(dz.median())
        if z.shape[1] > 1:
            fc.bias[0] -= m / 2
            fc.bias[1] += m / 2
        else:
            fc.bias[0] -= m
    return model


def gather_host_patches(slide, coords_xy, ps, pitch_px):
    """(n, ps, pitch_px, 4) uint8 host stack of read_region windows (zero padded outside the slide)."""
    n = len(coords_xy)
    out = torch.zeros((n, ps, pitch_px, 4), dtype=torch.uint8)
    W, H = slide.width, slide.height
    for s in range(0, n, 512):
        blk = coords_xy[s:s + 512]
        dev = torch.zeros((len(blk), ps, pitch_px, 4), dtype=torch.uint8, device=slide.rgba.device)
        for i, (x, y) in enumerate(blk.tolist()):
            x0, y0, x1, y1 = max(x, 0), max(y, 0), min(x + ps, W), min(y + ps, H)
            if x1 > x0 and y1 > y0:
                dev[i, y0 - y:y1 - y, x0 - x:x1 - x] = slide.rgba[y0:y1, x0:x1]
        out[s:s + len(blk)] = dev.cpu()
    return out


def oracle_preprocess(rgba_stack, ps, S, mean, std, pool):
    from oracle import preprocess

    fn = lambda p: preprocess.preprocess_lib(np.ascontiguousarray(p[:, :ps]), S, mean, std)  # noqa: E731
    return torch.stack(list(pool.map(fn, rgba_stack.numpy())), 0)


def run_parity(arch="resnet34", K=2, n_patches=10240, batch=1024, slide_wh=(32768, 24576), mpp=0.5,
               oracle_mode="cpu", anchor=512, device=0, seed=0, chunk=256, log=None):
    from oracle import forward as oforward
    from wsinsight_b200.modellib.engine import B200Network
    from wsinsight_b200.modellib.run_inference import HostPatchClassifier, classify_patches_device
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from wsinsight_b200.patchlib import segment_and_patch_slides_device
    from wsinsight_b200.wsi import DeviceSlide

    log = log or (lambda *a: None)
    t_all = time.perf_counter()
    torch.cuda.set_device(device)
    dev = torch.device("cuda", device)
    model, cfg = make_model(arch, K, seed)
    tfa = cfg["transform"][2]["arguments"]
    S = int(cfg["transform"][0]["arguments"]["size"])
    slide = DeviceSlide.synthetic(slide_wh[0], slide_wh[1], mpp=mpp, seed=seed, device=dev)
    res = segment_and_patch_slides_device([slide], cfg["patch_size_pixels"], cfg["spacing_um_px"])[0]
    ps = res.patch_size
    n_all = int(res.coords.shape[0])
    assert n_all >= n_patches, f"slide has only {n_all} in-tissue patches"
    idx = torch.linspace(0, n_all - 1, n_patches, device=dev).long()
    coords = res.coords[idx].contiguous()
    cxy = coords.cpu().numpy().astype(np.int64)
    xywh = np.concatenate([cxy, np.full_like(cxy, ps)], 1)
    pitch_px = (ps + 3) // 4 * 4
    host_rgba = gather_host_patches(slide, cxy, ps, pitch_px)
    pool = ThreadPoolExecutor(max_workers=8)

    # ---- classifier centred on the workload (fp32, CPU) ----
    x0 = oracle_preprocess(host_rgba[:256], ps, S, tfa["mean"], tfa["std"], pool)
    center_classifier(model, x0)
    ts = torch.jit.freeze(torch.jit.script(model))

    # ---- oracle + library comparators, chunked ----
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    m_fp32_gpu = copy.deepcopy(model).to(dev)
    m_bf16_gpu = copy.deepcopy(model).to(dev).to(torch.bfloat16).to(memory_format=torch.channels_last)
    m_wq = copy.deepcopy(model)
    with torch.no_grad():
        for p_ in m_wq.parameters():
            if p_.dim() == 4:
                p_.copy_(p_.to(torch.bfloat16).float())
    m_wq = m_wq.to(dev)
    want_cpu, want_gpu32, lib_bf16, wq_only = [], [], [], []
    n_cpu = n_patches if oracle_mode == "cpu" else min(anchor, n_patches)
    t0 = time.perf_counter()
    with torch.no_grad():
        for s in range(0, n_patches, chunk):
            x = oracle_preprocess(host_rgba[s:s + chunk], ps, S, tfa["mean"], tfa["std"], pool)
            if s < n_cpu:
                want_cpu.append(oforward.head(ts(x)))
            xd = x.to(dev)
            want_gpu32.append(oforward.head(m_fp32_gpu(xd)).cpu())
            wq_only.append(oforward.head(m_wq(xd)).cpu())
            xb = xd.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
            lib_bf16.append(oforward.head(m_bf16_gpu(xb).float()).cpu())
    t_oracle = time.perf_counter() - t0
    want_cpu, want_gpu32 = torch.cat(want_cpu), torch.cat(want_gpu32)
    lib_bf16, wq_only = torch.cat(lib_bf16), torch.cat(wq_only)
    anchor_diff = float((want_gpu32[:n_cpu] - want_cpu).abs().max())
    want = want_cpu if oracle_mode == "cpu" else want_gpu32
    log(f"oracle {oracle_mode}: {n_patches} patches in {t_oracle:.1f} s; cuDNN-fp32 vs CPU-fp32 on {n_cpu}: {anchor_diff:.2e}")

    # ---- product paths ----
    net = B200Network.from_torchscript(ts, K, input_size=S, max_batch=batch, device=dev)
    transform = make_compose_from_transform_config(cfg["transform"])
    p_res = classify_patches_device(slide, xywh, net, transform, batch, device_coords=coords).cpu()
    hp = HostPatchClassifier(net, ps, transform, batch, depth=2)
    p_host = torch.zeros((n_patches, K), dtype=torch.float32).pin_memory()
    pinned = host_rgba.pin_memory()
    for s in range(0, n_patches, batch):
        hp.submit(pinned[s:s + batch], p_host[s:s + batch])
    hp.sync()
    hp.close()
    p_host = p_host.clone()
    if K == 1:
        p_res, p_host = p_res.reshape(-1), p_host.reshape(-1)
    net.close()

    pf = _p_first(want)
    out = {
        "arch": arch, "num_classes": K, "n_patches": n_patches, "patch_px": ps, "input_size": S,
        "slide": f"{slide_wh[0]}x{slide_wh[1]}@{mpp}", "in_tissue_patches": n_all,
        "weights": ("BN-calibrated random-init torchvision resnet34" if arch == "resnet34"
                    else "He-init Cadene-layout InceptionV4 without BatchNorm") +
                   "; classifier scaled to decision-logit std 1.5 and centred on the workload median",
        "oracle": ("CPU fp32 TorchScript forward + PIL/torchvision preprocess" if oracle_mode == "cpu" else
                   f"cuDNN fp32 (TF32 off) forward on the oracle's PIL/torchvision tensors, anchored by the CPU fp32 "
                   f"forward on the first {n_cpu} patches"),
        "anchor_max_abs_diff_cudnn_fp32_vs_cpu_fp32": anchor_diff, "anchor_patches": n_cpu,
        "spread_frac_p_in_0.05_0.95": float(((pf > 0.05) & (pf < 0.95)).double().mean()),
        "p_quantiles": [float(v) for v in torch.quantile(pf.double(), torch.tensor([0, .1, .25, .5, .75, .9, 1.0],
                                                                                    dtype=torch.float64))],
        "b200_resident": diff_stats(p_res, want), "b200_hostpipe": diff_stats(p_host, want),
        "hostpipe_vs_resident_max_abs": float((p_res - p_host).abs().max()),
        "cudnn_bf16_channels_last": diff_stats(lib_bf16, want),
        "fp32_with_bf16_rounded_conv_weights_only": diff_stats(wq_only, want),
        "north_star_tolerance": {"max_abs_dp": 2e-3, "argmax_agreement": 0.999},
        "seconds": time.perf_counter() - t_all, "oracle_seconds": t_oracle,
    }
    return out
