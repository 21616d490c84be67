"""The reference's command line for the hot path (wsinsight/cli/run.py:163-691, cli/patch.py:603-615)."""

import json
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest
from click.testing import CliRunner

ROOT = Path(__file__).resolve().parents[1]


def _bundle(tmp_path, calibrated=False):
    import torch
    from oracle import forward as oforward

    d = tmp_path / "model"
    d.mkdir(exist_ok=True)
    m = oforward.make_resnet34_calibrated(2, 0) if calibrated else oforward.make_resnet34(2, 0)
    oforward.export_torchscript(m, d / "torchscript_model.pt", frozen=True)
    (d / "config.json").write_text(json.dumps(oforward.BREAST_RESNET34_CONFIG))
    return m, d / "config.json", d / "torchscript_model.pt"


def test_cli_validation_rules(tmp_path):
    """Option names / defaults parse and the reference's XOR rules raise UsageError (exit code 2) before any GPU work."""
    from wsinsight_b200.cli import cli

    r = CliRunner()
    wsi = tmp_path / "slides"
    wsi.mkdir()
    out = tmp_path / "res"
    base = ["-i", str(wsi), "-o", str(out)]
    res = r.invoke(cli, ["run"] + base)
    assert res.exit_code == 2 and "one of --model or (--config and --model-path)" in res.output
    cfg = tmp_path / "config.json"
    from oracle.forward import BREAST_RESNET34_CONFIG
    cfg.write_text(json.dumps(BREAST_RESNET34_CONFIG))
    res = r.invoke(cli, ["run"] + base + ["--config", str(cfg)])
    assert res.exit_code == 2 and "--config and --model-path must both be set if one is set." in res.output
    pt = tmp_path / "m.pt"
    pt.write_bytes(b"x")
    res = r.invoke(cli, ["patch"] + base + ["--config", str(cfg), "--model-path", str(pt), "--model", "breast-tumor-resnet34.tcga-brca"])
    assert res.exit_code == 2 and "mutually exclusive with --model" in res.output
    res = r.invoke(cli, ["infer"] + base + ["--config", str(cfg), "--model-path", str(pt), "--omecsv"])
    assert res.exit_code == 2 and "--omecsv" in res.output
    res = r.invoke(cli, ["run"] + base + ["--config", str(cfg), "--model-path", str(pt), "--synthetic", "bogus"])
    assert res.exit_code == 2 and "WxH@mpp" in res.output
    res = r.invoke(cli, ["run"] + base + ["--config", str(cfg), "--model-path", str(pt), "--seg-median-filter-size", "2"])
    assert res.exit_code == 2                                     # IntRange(min=3) as upstream
    res = r.invoke(cli, ["run"] + base + ["--config", str(cfg), "--model-path", str(pt), "--patch-overlap-ratio", "1.0"])
    assert res.exit_code == 2                                     # (-inf, 1)
    res = r.invoke(cli, ["run", "--help"])
    for opt in ("--wsi-dir", "--results-dir", "--model-path", "--batch-size", "--num-workers", "--geojson",
                "--seg-thumbsize", "--seg-median-filter-size", "--seg-binary-threshold", "--seg-closing-kernel-size",
                "--seg-min-object-size-um2", "--seg-min-hole-size-um2", "--patch-overlap-ratio", "--patch-size-um",
                "--patch-size-px"):
        assert opt in res.output, opt


def test_cli_helpers():
    from wsinsight_b200.cli import _overlap, _parse_synth_name, _synthetic_path, enumerate_slide_paths
    from wsinsight_b200.modellib.models import ModelConfiguration
    from oracle.forward import BREAST_RESNET34_CONFIG

    cfg = ModelConfiguration.from_dict(BREAST_RESNET34_CONFIG)
    assert _overlap(cfg, 0.0, 0.0, 0) == 0.0 and _overlap(cfg, 0.25, 0.0, 0) == 0.25
    assert _overlap(cfg, 0.0, 43.75, 0) == pytest.approx(0.5)       # 350 px * 0.25 um = 87.5 um
    assert _overlap(cfg, 0.0, 0.0, 175) == pytest.approx(0.5)
    p = _synthetic_path(Path("/x"), 4096, 3072, 0.25, 3)
    assert _parse_synth_name(p) == dict(width=4096, height=3072, mpp=0.25, seed=3)
    assert [q.name for q in enumerate_slide_paths(Path("/nonexistent-dir"), ("100x80@0.5", "64x64@1:7"))] == \
        ["synthetic_100x80_mpp0.5_s0.synth", "synthetic_64x64_mpp1_s7.synth"]


def test_mask_overlay_colours():
    """masks/<stem>.jpg content (reference io.py:146-166): outer contours cyan, hole contours yellow, 7 px wide."""
    import cv2 as cv
    from conftest import blob_mask, blob_thumbnail
    from wsinsight_b200.patchlib.pipeline import draw_contours_on_thumbnail

    m = blob_mask(300, 400, 1)
    th = blob_thumbnail(300, 400, 1)
    got = draw_contours_on_thumbnail(th, m)
    contours, hier = cv.findContours(m.astype("uint8") * 255, cv.RETR_CCOMP, cv.CHAIN_APPROX_SIMPLE)
    want = th.copy()
    ext = (hier[0, :, 3] < 0).tolist()
    cv.drawContours(want, [c for c, e in zip(contours, ext) if e], -1, (0, 255, 255), 7)
    cv.drawContours(want, [c for c, e in zip(contours, ext) if not e], -1, (255, 255, 0), 7)
    assert np.array_equal(got, want)
    assert (got == np.array([0, 255, 255])).all(-1).sum() > 100 and (got == np.array([255, 255, 0])).all(-1).sum() > 10
    assert np.array_equal(draw_contours_on_thumbnail(th, np.zeros_like(m)), th)


@pytest.mark.gpu
def test_cli_run_produces_reference_outputs(tmp_path):
    """`run` on a synthetic slide: patches/, masks/, model-outputs-csv/, model-outputs-geojson/, metadata JSONs; the CSV
    rows are the oracle's coordinates and the probabilities of the device path; a second run skips finished slides."""
    import pandas as pd
    import torch
    from PIL import Image

    from oracle import patchgrid, segment
    from wsinsight_b200.cli import cli
    from wsinsight_b200.wsi import DeviceSlide

    model, cfg, pt = _bundle(tmp_path, calibrated=True)
    wsi = tmp_path / "slides"
    wsi.mkdir()
    (wsi / "a.synth").write_text(json.dumps(dict(width=4096, height=3072, mpp=0.25, seed=1)))
    out = tmp_path / "res"
    args = ["run", "-i", str(wsi), "-o", str(out), "--config", str(cfg), "--model-path", str(pt), "-b", "16",
            "--geojson", "--seg-thumbsize", "512", "512", "--seg-min-object-size-um2", "6400",
            "--seg-min-hole-size-um2", "4900", "--synthetic", "3000x2500@0.5:2"]
    res = CliRunner().invoke(cli, args, catch_exceptions=False)
    assert res.exit_code == 0, res.output
    stems = ["a", "synthetic_3000x2500_mpp0.5_s2"]
    for st in stems:
        assert (out / "patches" / f"{st}.h5").exists() or (out / "patches" / f"{st}.h5.npz").exists()
        assert (out / "masks" / f"{st}.jpg").exists()
        assert (out / "model-outputs-csv" / f"{st}.csv").exists()
        assert (out / "model-outputs-geojson" / f"{st}.json").exists() or \
            (out / "model-outputs-geojson" / f"{st}.geojson").exists()
    assert list(out.glob("patch_metadata_*.json")) and list(out.glob("infer_metadata_*.json"))
    im = Image.open(out / "masks" / "a.jpg")
    assert max(im.size) <= 1024
    # CSV content vs the oracle's coordinates
    slide = DeviceSlide.synthetic(4096, 3072, 0.25, seed=1)
    thumb = slide.thumbnail_device((512, 512)).cpu().numpy()
    sc = patchgrid.slide_params((4096, 3072), (thumb.shape[1], thumb.shape[0]), 0.25, 350, 0.25, 6400, 4900)
    mask = segment.segment_tissue(thumb, 7, 7, 6, sc["min_object_size_px"], sc["min_hole_size_px"])
    cent, _, _ = patchgrid.grid_centroids(4096, 3072, sc["patch_size"], sc["half_patch_size"])
    keep = patchgrid.raster_rule(segment.outermost_restated(mask), cent, 4096, 3072)
    df = pd.read_csv(out / "model-outputs-csv" / "a.csv")
    assert list(df.columns) == ["minx", "miny", "width", "height", "prob_Other", "prob_Tumor"]
    assert np.array_equal(df[["minx", "miny"]].to_numpy(), cent[keep] - sc["half_patch_size"])
    assert (df["width"] == sc["patch_size"]).all() and np.allclose(df["prob_Other"] + df["prob_Tumor"], 1, atol=1e-5)
    # resume: nothing is recomputed, files untouched
    before = {p: p.stat().st_mtime_ns for p in (out / "model-outputs-csv").iterdir()}
    res = CliRunner().invoke(cli, args, catch_exceptions=False)
    assert res.exit_code == 0
    assert before == {p: p.stat().st_mtime_ns for p in (out / "model-outputs-csv").iterdir()}
    # patch + infer as separate commands give the same CSV bytes
    out2 = tmp_path / "res2"
    common = ["-i", str(wsi), "-o", str(out2), "--config", str(cfg), "--model-path", str(pt)]
    r1 = CliRunner().invoke(cli, ["patch"] + common + ["--seg-thumbsize", "512", "512", "--seg-min-object-size-um2", "6400",
                                                      "--seg-min-hole-size-um2", "4900"], catch_exceptions=False)
    r2 = CliRunner().invoke(cli, ["infer"] + common + ["-b", "16"], catch_exceptions=False)
    assert r1.exit_code == 0 and r2.exit_code == 0, r1.output + r2.output
    assert (out2 / "model-outputs-csv" / "a.csv").read_bytes() == (out / "model-outputs-csv" / "a.csv").read_bytes()


@pytest.mark.gpu
def test_cli_two_ranks_same_csv_as_one(tmp_path):
    """The multi-GPU product path (replaces nn.DataParallel, reference run_inference.py:173-174): two ranks under torchrun
    shard ONE slide's patch list, gather over NCCL, rank 0 writes the CSV -- byte-identical to the single-GPU CSV."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _, cfg, pt = _bundle(tmp_path, calibrated=True)
    wsi = tmp_path / "slides"
    wsi.mkdir()
    (wsi / "a.synth").write_text(json.dumps(dict(width=6000, height=5000, mpp=0.5, seed=4)))
    common = ["-i", str(wsi), "--config", str(cfg), "--model-path", str(pt), "-b", "64", "--seg-thumbsize", "512", "512",
              "--seg-min-object-size-um2", "6400", "--seg-min-hole-size-um2", "4900"]
    env = dict(os.environ, PYTHONPATH=str(ROOT))
    one = subprocess.run([sys.executable, "-m", "wsinsight_b200", "run", "-o", str(tmp_path / "one")] + common,
                         env=env, capture_output=True, text=True, cwd=ROOT)
    assert one.returncode == 0, one.stdout + one.stderr
    two = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", "-m", "wsinsight_b200", "run",
                          "-o", str(tmp_path / "two")] + common, env=env, capture_output=True, text=True, cwd=ROOT)
    assert two.returncode == 0, two.stdout + two.stderr
    a = (tmp_path / "one" / "model-outputs-csv" / "a.csv").read_bytes()
    b = (tmp_path / "two" / "model-outputs-csv" / "a.csv").read_bytes()
    assert len(a) > 1000 and a == b
    # the same with the peer-store result path: the tail kernel writes into rank 0's buffer over NVLink, no gather
    peer = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                           "--master-addr", "127.0.0.1", "--master-port", "29534", "-m", "wsinsight_b200", "run",
                           "-o", str(tmp_path / "peer")] + common, env=dict(env, WSI_B200_RESULT_PATH="peer"),
                          capture_output=True, text=True, cwd=ROOT)
    assert peer.returncode == 0, peer.stdout + peer.stderr
    assert (tmp_path / "peer" / "model-outputs-csv" / "a.csv").read_bytes() == a


@pytest.mark.gpu
def test_cli_purple_slide_matches_reference_golden_csv(tmp_path, golden_purple):
    """Mirror of the reference's integration test (tests/test_all.py:413-478 `test_cli_run_with_local_model`, :559-612
    `test_patch_cli`): `run` on the uniform purple 4096 x 4096 slide at 0.25 um/px.  The slide is a flat PNG read through
    PIL (decoders are out of scope); the coordinate columns must equal the reference's golden purple.csv row for row
    (the probability columns need the real zoo weights and stay unpinned), and the 'patch' expectations of the
    reference's grid formula hold for its five (patch size, spacing) pairs."""
    import pandas as pd
    from PIL import Image

    from oracle import forward as oforward
    from wsinsight_b200.cli import cli
    from wsinsight_b200.patchlib.io import load_patch_file

    wsi = tmp_path / "slides"
    wsi.mkdir()
    g = golden_purple["breast-tumor-resnet34.tcga-brca"]
    Image.new("RGB", (4096, 4096), tuple(g["rgb"])).save(wsi / "purple.png")
    _, cfg, pt = _bundle(tmp_path)
    out = tmp_path / "inference"
    res = CliRunner().invoke(cli, ["run", "--wsi-dir", str(wsi), "--results-dir", str(out), "--model-path", str(pt),
                                   "--config", str(cfg), "--slide-mpp", "0.25"], catch_exceptions=False)
    assert res.exit_code == 0, res.output
    df = pd.read_csv(out / "model-outputs-csv" / "purple.csv")
    assert list(df.columns) == g["columns"]
    assert df[["minx", "miny"]].to_numpy().tolist() == g["coords"]
    assert (df["width"] == g["patch_size"]).all() and (df["height"] == g["patch_size"]).all()
    assert (out / "masks" / "purple.jpg").exists()
    # 'patch' alone for the reference's five (patch_size, spacing) pairs: its expected grid (test_all.py:590-603)
    for k, (psz, spacing) in enumerate([(256, 0.25), (256, 0.50), (350, 0.25), (100, 0.3), (100, 0.5)]):
        c = dict(oforward.BREAST_RESNET34_CONFIG, patch_size_pixels=psz, spacing_um_px=spacing)
        cp = tmp_path / f"cfg{k}.json"
        cp.write_text(json.dumps(c))
        sd = tmp_path / f"savedir{k}"
        r = CliRunner().invoke(cli, ["patch", "--wsi-dir", str(wsi), "--results-dir", str(sd), "--model-path", str(pt),
                                     "--config", str(cp), "--slide-mpp", "0.25"], catch_exceptions=False)
        assert r.exit_code == 0, r.output
        eps = round(psz * spacing / 0.25)
        want = [[x, y] for x in range(0, 4096, eps) for y in range(0, 4096, eps)
                if x + eps // 2 <= 4096 and y + eps // 2 <= 4096]
        assert len(want) == round(4096 / eps) ** 2
        d = load_patch_file(sd / "patches" / "purple.h5")
        assert int(d["patch_size"]) == eps
        assert np.asarray(d["coords"]).tolist() == want, (psz, spacing)
