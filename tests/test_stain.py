"""N4: Macenko stain normalisation (reference run_inference.py:232-264, data.py:293-300).  HistomicsTK is not importable:
oracle/stain.py restates its call chain -- parity with the library itself is UNPINNED; these tests pin the product's host
estimation and the CUDA kernel to that restatement."""

import numpy as np
import pytest

from oracle import stain as ost


def _he_image(h, w, seed):
    rng = np.random.default_rng(seed)
    W = np.array([[0.65, 0.70, 0.29], [0.07, 0.99, 0.11]]).T
    c = rng.gamma(2.0, 0.6, (2, h * w))
    sda = W @ c * 60
    rgb = np.clip(255 * np.exp(-sda * np.log(255) / 255) + rng.normal(0, 2, sda.shape), 0, 255)
    return rgb.T.reshape(h, w, 3).astype(np.uint8)


def test_host_estimation_matches_restated_library_chain():
    from wsinsight_b200.modellib import stain as pst

    rgb = _he_image(100, 200, 0)
    a = ost.rgb_separate_stains_macenko_pca(rgb.astype(np.float32) + ost.EPSILON, ost.I_0)
    b = pst.macenko_stain_matrix(rgb.astype(np.float32) + np.float32(pst.EPSILON), pst.I_0)
    assert np.abs(a - b).max() < 1e-6
    assert np.allclose(np.linalg.norm(a, axis=0), 1.0)
    assert np.allclose(pst.default_target_matrix(), ost.default_target_matrix())
    # normalising with source == target (unit columns) is the identity on the stain densities
    M = pst.normalization_matrix(a, a)
    assert np.abs(M - np.eye(3)).max() < 1e-9


def test_oracle_chain_shapes():
    rgb = _he_image(40, 50, 1)
    W = ost.rgb_separate_stains_macenko_pca(rgb.astype(np.float32) + ost.EPSILON, ost.I_0)
    out = ost.normalize_patch(rgb, W, ost.default_target_matrix())
    assert out.shape == rgb.shape and out.dtype == np.uint8
    white = np.full((4, 4, 3), 255, np.uint8)
    assert ost.normalize_patch(white, W, ost.default_target_matrix()).min() >= 254      # background stays white


@pytest.mark.gpu
def test_kernel_matches_restated_chain():
    import torch

    from wsinsight_b200.modellib import stain as pst

    rgb = _he_image(175, 350, 2)
    rgb[:5] = np.random.default_rng(3).integers(0, 256, rgb[:5].shape)       # arbitrary colours incl. 0 and 255
    rgb[5, :3] = [[0, 0, 0], [255, 255, 255], [0, 255, 0]]
    W = ost.rgb_separate_stains_macenko_pca(rgb.astype(np.float32) + ost.EPSILON, ost.I_0)
    want = ost.normalize_patch(rgb, W, ost.default_target_matrix())
    got = pst.normalize_patches_device(torch.from_numpy(rgb).cuda(), W, pst.default_target_matrix()).cpu().numpy()
    d = np.abs(got.astype(int) - want.astype(int))
    # float64 log / pow of the device vs libm: a truncation may flip on a value within 1e-12 of an integer
    assert d.max() <= 1 and (d > 0).mean() < 1e-4, (d.max(), (d > 0).mean())


@pytest.mark.gpu
def test_run_inference_with_stain_normalization(tmp_path):
    """End to end: run_inference(stain_normalization=True) == the oracle chain (normalise each patch, then the same
    classification path from images) on a small synthetic slide."""
    import json

    import pandas as pd
    import torch

    from oracle import forward as oforward
    from wsinsight_b200.modellib.engine import B200Network
    from wsinsight_b200.modellib.models import LocalModelTorchScript, ModelConfiguration
    from wsinsight_b200.modellib.run_inference import classify_patches_from_images, run_inference
    from wsinsight_b200.modellib.stain import default_target_matrix, estimate_slide_stain_matrix
    from wsinsight_b200.modellib.transforms import make_compose_from_transform_config
    from wsinsight_b200.patchlib import segment_and_patch_directory_of_slides
    from wsinsight_b200.patchlib.io import extract_patches_from_slide
    from wsinsight_b200.wsi import DeviceSlide

    model = oforward.make_resnet34_calibrated(2, 0)
    d = tmp_path / "m"
    d.mkdir()
    oforward.export_torchscript(model, d / "torchscript_model.pt")
    cfg = ModelConfiguration.from_dict(oforward.BREAST_RESNET34_CONFIG)
    slide = DeviceSlide.synthetic(3000, 2600, 0.5, seed=5)
    sp = str(tmp_path / "s.synth")
    res = tmp_path / "res"
    res.mkdir()
    segment_and_patch_directory_of_slides(None, [sp], res, patch_size_px=cfg.patch_size_pixels,
                                          patch_spacing_um_px=cfg.spacing_um_px, slides={sp: slide}, thumbsize=(512, 512),
                                          min_object_size_um2=6400, min_hole_size_um2=4900)
    fp, fi = run_inference(None, [sp], res, model_info=LocalModelTorchScript(cfg, str(d / "torchscript_model.pt")),
                           batch_size=16, stain_normalization=True, slides={sp: slide})
    assert fp == [] and fi == []
    df = pd.read_csv(res / "model-outputs-csv" / "s.csv")
    xywh = df[["minx", "miny", "width", "height"]].to_numpy()
    assert len(xywh) > 20
    W = estimate_slide_stain_matrix(slide, xywh, 256, seed=0)
    px = extract_patches_from_slide(slide, xywh[:, :2], int(xywh[0, 2]), as_numpy=True)
    norm = np.stack([ost.normalize_patch(p, W, ost.default_target_matrix()) for p in px[:24]])
    net = B200Network.from_torchscript(str(d / "torchscript_model.pt"), 2, input_size=224, max_batch=16)
    tf = make_compose_from_transform_config(oforward.BREAST_RESNET34_CONFIG["transform"])
    want = classify_patches_from_images(norm, net, tf, 16).cpu().numpy()
    got = df[["prob_Other", "prob_Tumor"]].to_numpy()[:24]
    assert np.abs(got - want).max() < 5e-3          # identical pixels up to rare 1-LSB truncation flips
    # and normalisation does change the result
    plain = classify_patches_from_images(px[:24], net, tf, 16).cpu().numpy()
    assert np.abs(plain - want).max() > 1e-3
