"""``python -m wsinsight_b200 run|patch|infer`` -- the reference's command line for the hot path.

Mirror of ``wsinsight/cli/run.py:163-691`` (``run`` = ``patch`` then ``infer`` with the same arguments),
``wsinsight/cli/patch.py`` (model resolution rules :603-615, overlap from ``--patch-size-um/px`` :833-851, metadata
JSON :889-894) and ``wsinsight/cli/infer.py`` (:1060-1120 inference + GeoJSON export, :1296-1301 metadata).  Option
names and defaults are the reference's.  Options that select branches outside the B200 hot path (QuPath inputs,
HistoQC, object models, OME-CSV, H-Plot, CME) are accepted so existing command lines parse, and raise a
``UsageError`` when used.

Slide decoders are out of scope (tiffslide / openslide are not in the image), so ``--wsi-dir`` may contain

* ``*.synth`` JSON stubs ``{"width": W, "height": H, "mpp": M, "seed": S}`` -- procedural slides generated on the
  device (``--synthetic WxH@mpp[:seed]`` adds such slides without touching the directory), and
* flat images PIL can read (``.png .jpg .jpeg .tif .tiff .bmp``) with the resolution given by ``--slide-mpp``.

Multi-GPU: launched under ``torchrun`` (one process per GPU) every rank runs the same command; each slide's patch
list is sharded across the ranks and rank 0 writes the outputs (``modellib/run_inference.py``).
"""

from __future__ import annotations

import json
import logging
import os
import re
import sys
from datetime import datetime
from pathlib import Path
from typing import Dict, List, Optional

import click

IMAGE_SUFFIXES = {".png", ".jpg", ".jpeg", ".tif", ".tiff", ".bmp"}
_SYNTH_RE = re.compile(r"^(\d+)x(\d+)@([0-9.]+)(?::(\d+))?$")


def _num_cpus() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _synthetic_path(wsi_dir: Path, w: int, h: int, mpp: float, seed: int) -> Path:
    return Path(wsi_dir) / f"synthetic_{w}x{h}_mpp{mpp:g}_s{seed}.synth"


def _parse_synth_name(path: Path) -> Optional[dict]:
    m = re.match(r"^synthetic_(\d+)x(\d+)_mpp([0-9.]+)_s(\d+)$", Path(path).stem)
    if not m:
        return None
    return dict(width=int(m.group(1)), height=int(m.group(2)), mpp=float(m.group(3)), seed=int(m.group(4)))


def enumerate_slide_paths(wsi_dir: Path, synthetic: tuple = ()) -> List[Path]:
    """Files of the slide directory (reference ``_enumerate_slide_paths``) plus ``--synthetic`` slides."""
    paths = sorted(p for p in Path(wsi_dir).iterdir() if p.is_file()) if Path(wsi_dir).exists() else []
    for spec in synthetic:
        m = _SYNTH_RE.match(spec)
        if not m:
            raise click.UsageError(f"--synthetic expects WxH@mpp[:seed], got {spec!r}")
        paths.append(_synthetic_path(wsi_dir, int(m.group(1)), int(m.group(2)), float(m.group(3)), int(m.group(4) or 0)))
    return paths


def open_slide(path: Path, slide_mpp: Optional[float] = None):
    """Path -> :class:`~wsinsight_b200.wsi.DeviceSlide` on the current device."""
    from .wsi import DeviceSlide

    path = Path(path)
    if path.suffix == ".synth":
        spec = json.loads(path.read_text()) if path.exists() else _parse_synth_name(path)
        if spec is None:
            raise ValueError(f"cannot derive a synthetic slide from {path}")
        return DeviceSlide.synthetic(int(spec["width"]), int(spec["height"]), float(spec["mpp"]),
                                     seed=int(spec.get("seed", 0)), name=path.stem)
    if path.suffix.lower() in IMAGE_SUFFIXES:
        import numpy as np
        from PIL import Image

        Image.MAX_IMAGE_PIXELS = None
        if slide_mpp is None:
            raise click.UsageError(f"--slide-mpp is required to read flat image {path.name}")
        with Image.open(path) as im:
            arr = np.asarray(im.convert("RGBA"))
        return DeviceSlide.from_numpy(arr, slide_mpp, name=path.stem)
    raise ValueError(f"no decoder for {path.name}: slide decoders (tiffslide/openslide) are outside the B200 hot path")


def _resolve_model(model_name, config, model_path, qupath_detection_dir, qupath_geojson_detection_dir,
                   qupath_geojson_annotation_dir):
    """Reference rules (cli/patch.py:603-615, cli/infer.py:786-815)."""
    from .modellib.models import LocalModelTorchScript, ModelConfiguration, get_registered_model

    qp = (qupath_detection_dir, qupath_geojson_detection_dir, qupath_geojson_annotation_dir)
    if model_name is None and config is None and model_path is None and all(q is None for q in qp):
        raise click.UsageError(
            "one of --model or (--config and --model-path) or --qupath_detection_dir or --qupath_geojson_detection_dir "
            "or --qupath_geojson_annotation_dir is required.")
    if (config is not None or model_path is not None) and model_name is not None:
        raise click.UsageError("--config and --model-path are mutually exclusive with --model.")
    if (config is not None) ^ (model_path is not None):
        raise click.UsageError("--config and --model-path must both be set if one is set.")
    if any(q is not None for q in qp):
        raise click.UsageError("QuPath detection / GeoJSON pseudo-models are outside the B200 hot path")
    if model_name is not None:
        get_registered_model(model_name)          # raises: the zoo registry needs network access
    return LocalModelTorchScript(ModelConfiguration.from_json(config), str(model_path))


def _overlap(cfg, patch_overlap_ratio, patch_size_um, patch_size_px):
    """cli/patch.py:833-851."""
    if patch_overlap_ratio != 0.0:
        return patch_overlap_ratio
    if patch_size_um != 0.0:
        if patch_size_um > cfg.patch_size_pixels * cfg.spacing_um_px:
            raise click.ClickException("--patch-size-um has to be smaller than patch size")
        return 1.0 - patch_size_um / (cfg.patch_size_pixels * cfg.spacing_um_px)
    if patch_size_px != 0:
        if patch_size_px > cfg.patch_size_pixels:
            raise click.ClickException("--patch-size-px must not be larger than patch size")
        return 1.0 - float(patch_size_px) / float(cfg.patch_size_pixels)
    return 0.0


def _run_metadata(model_obj) -> dict:
    """What ``_get_info_for_save`` records (cli/patch.py:122-200), for this implementation."""
    import torch

    from . import _cabi

    return {
        "model_weights": {"path": str(model_obj.model_path)},
        "config": {"architecture": model_obj.config.architecture, "num_classes": model_obj.config.num_classes,
                   "class_names": model_obj.config.class_names,
                   "patch_size_pixels": model_obj.config.patch_size_pixels,
                   "spacing_um_px": model_obj.config.spacing_um_px},
        "runtime": {"implementation": "wsinsight_b200", "library": _cabi.load().wsi_version().decode(),
                    "python_executable": sys.executable, "python_version": sys.version,
                    "torch_version": torch.__version__, "cuda_device": torch.cuda.get_device_name(),
                    "timestamp": datetime.now().astimezone().isoformat(), "cli": sys.argv},
    }


def _unsupported(**flags):
    bad = [k for k, v in flags.items() if v]
    if bad:
        raise click.UsageError("outside the B200 hot path (see DESIGN.md section 2): " +
                               ", ".join("--" + b.replace("_", "-") for b in bad))


def _rank_world():
    from .modellib.run_inference import dist_rank_world

    return dist_rank_world()


def _init_distributed():
    """Under torchrun (WORLD_SIZE > 1) bind this process to its GPU and join the NCCL job."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    import torch

    if world > 1:
        import torch.distributed as dist

        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))


# ------------------------------------------------------------------------------------------------
# shared options
# ------------------------------------------------------------------------------------------------
def _common_options(f):
    opts = [
        click.option("-i", "--wsi-dir", type=click.Path(path_type=Path), required=True,
                     help="Directory containing whole slide images. This directory can *only* contain whole slide images."),
        click.option("-o", "--results-dir", type=click.Path(path_type=Path), required=True,
                     help="Directory to store results. If directory exists, will skip whole slides for which outputs exist."),
        click.option("-r", "--references-dir", type=click.Path(path_type=Path), default=None),
        click.option("--qupath-detection-dir", type=click.Path(path_type=Path), default=None),
        click.option("--qupath-geojson-detection-dir", type=click.Path(path_type=Path), default=None),
        click.option("--qupath-geojson-annotation-dir", type=click.Path(path_type=Path), default=None),
        click.option("--qupath-detection-patch-size", default=56, type=click.IntRange(min=1)),
        click.option("--qupath-annotation-patch-size", default=224, type=click.IntRange(min=1)),
        click.option("--qupath-spacing-um-px", default=0.5, type=click.FloatRange(min=0)),
        click.option("--qupath-name-as-class", is_flag=True, default=False, show_default=True),
        click.option("-m", "--model", "model_name", default=None,
                     help="Name of the model to use from WSInsight Model Zoo. Mutually exclusive with --config."),
        click.option("-c", "--config", type=click.Path(exists=True, dir_okay=False, path_type=Path),
                     help="Path to configuration for the trained model. Mutually exclusive with --model"),
        click.option("-p", "--model-path", type=click.Path(exists=True, dir_okay=False, path_type=Path),
                     help="Path to the pretrained model. Use only when --config is passed."),
        click.option("--synthetic", multiple=True, metavar="WxH@mpp[:seed]",
                     help="(extension) add a procedural slide generated on the device; repeatable."),
        click.option("--slide-mpp", type=click.FloatRange(min=0, min_open=True), default=None,
                     help="(extension) microns per pixel of flat image files in --wsi-dir."),
    ]
    for o in reversed(opts):
        f = o(f)
    return f


def _patch_options(f):
    opts = [
        click.option("--cache-image-patches", is_flag=True, default=False, show_default=True),
        click.option("--histoqc-dir", type=click.Path(path_type=Path), default=None),
        click.option("--seg-thumbsize", default=(2048, 2048), type=(int, int)),
        click.option("--seg-median-filter-size", default=7, type=click.IntRange(min=3),
                     help="The kernel size for median filtering. Must be greater than 1 and odd."),
        click.option("--seg-binary-threshold", default=7, type=click.IntRange(min=1),
                     help="The threshold for image binarization."),
        click.option("--seg-otsu", is_flag=True, default=False,
                     help="(extension) choose the binarization threshold per slide by Otsu's method."),
        click.option("--seg-closing-kernel-size", default=6, type=click.IntRange(min=1)),
        click.option("--seg-closing-convention", default="scipy_raw",
                     type=click.Choice(["scipy_raw", "front_padded", "mirrored_erosion"]),
                     help="(extension) footprint-origin convention of the even-sized closing."),
        click.option("--seg-min-object-size-um2", default=200 ** 2, type=click.FloatRange(min=0)),
        click.option("--seg-min-hole-size-um2", default=190 ** 2, type=click.FloatRange(min=0)),
        click.option("--patch-overlap-ratio", default=0.0, type=click.FloatRange(min=None, max=1, max_open=True)),
        click.option("--patch-size-um", default=0.0, type=click.FloatRange(min=0.0)),
        click.option("--patch-size-px", default=0, type=click.FloatRange(min=0)),
    ]
    for o in reversed(opts):
        f = o(f)
    return f


def _infer_options(f):
    opts = [
        click.option("-b", "--batch-size", type=click.IntRange(min=1), default=32, show_default=True,
                     help="Batch size during model inference."),
        click.option("-n", "--num-workers", default=min(_num_cpus(), 8), show_default=True, type=click.IntRange(min=0),
                     help="Accepted for compatibility: there are no DataLoader workers on this path."),
        click.option("--qupath", is_flag=True, default=False, show_default=True),
        click.option("--geojson", is_flag=True, default=False, show_default=True,
                     help="Create a GeoJSON directory containing the inference results"),
        click.option("--omecsv", is_flag=True, default=False, show_default=True),
        click.option("--hplot", is_flag=True, default=False, show_default=True),
        click.option("--cme-cellular", is_flag=True, default=False, show_default=True),
        click.option("--cme-annotation", is_flag=True, default=False, show_default=True),
    ]
    for o in reversed(opts):
        f = o(f)
    return f


# ------------------------------------------------------------------------------------------------
# stages
# ------------------------------------------------------------------------------------------------
def do_patch(*, wsi_dir, results_dir, model_obj, slide_paths, slides, seg_thumbsize, seg_median_filter_size,
             seg_binary_threshold, seg_otsu, seg_closing_kernel_size, seg_closing_convention,
             seg_min_object_size_um2, seg_min_hole_size_um2, overlap):
    from .patchlib import segment_and_patch_directory_of_slides

    rank, _ = _rank_world()
    if not Path(wsi_dir).exists() and not slide_paths:
        raise FileNotFoundError(f"Whole slide image directory not found: {wsi_dir}")
    if not slide_paths:
        raise FileNotFoundError(f"no files exist in the slide directory: {wsi_dir}")
    click.secho("\nFinding patch coordinates...\n", fg="green")
    if rank == 0:
        Path(results_dir).mkdir(parents=True, exist_ok=True)
        segment_and_patch_directory_of_slides(
            wsi_dir=wsi_dir, slide_paths=slide_paths, save_dir=results_dir,
            patch_size_px=model_obj.config.patch_size_pixels, patch_spacing_um_px=model_obj.config.spacing_um_px,
            slides=slides, thumbsize=tuple(seg_thumbsize), median_filter_size=seg_median_filter_size,
            binary_threshold="otsu" if seg_otsu else seg_binary_threshold, closing_kernel_size=seg_closing_kernel_size,
            closing_convention=seg_closing_convention, min_object_size_um2=seg_min_object_size_um2,
            min_hole_size_um2=seg_min_hole_size_um2, overlap=overlap)
        if not (Path(results_dir) / "patches").exists():
            raise click.ClickException(
                "No patches were created. Please see the logs above and check for errors."
                " It is possible that no tissue was detected in the slides. If that is the"
                " case, please try to use different --seg-* parameters, which will change"
                " how the segmentation is done. For example, a lower binary threshold may be set.")
        ts = datetime.now().astimezone().strftime("%Y%m%dT%H%M%S")
        out = Path(results_dir) / f"patch_metadata_{ts}.json"
        click.echo(f"\nSaving metadata about run to {out}\n")
        out.write_text(json.dumps(_run_metadata(model_obj), indent=2))
    _barrier()
    click.secho("\nWSInsight-patch tasks are all finished.\n", fg="green")


def _barrier():
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def do_infer(*, wsi_dir, results_dir, model_obj, slide_paths, slides, batch_size, num_workers, geojson, overlap):
    from .modellib.run_inference import run_inference
    from .write_geojson import write_geojsons

    rank, _ = _rank_world()
    if not (Path(results_dir) / "patches").exists():
        raise click.ClickException(
            "No patches were created. Please see the logs above and check for errors. "
            "It is possible that no tissue was detected in the slides. If that is the case,"
            " please try to use different --seg-* parameters, which will change how the"
            " segmentation is done. For example, a lower binary threshold may be set.")
    if not slide_paths:
        raise FileNotFoundError(f"no files exist in the slide directory: {wsi_dir}")
    click.secho("\nRunning model inference.\n", fg="green")
    failed_patching, failed_inference = run_inference(
        wsi_dir=wsi_dir, slide_paths=slide_paths, results_dir=results_dir, model_info=model_obj,
        batch_size=batch_size, num_workers=num_workers, slides=slides,
        # the reference reads this switch from the model's config.json (cli/infer.py:845)
        stain_normalization=bool(model_obj.config.extras.get("stain_normalization")))
    if failed_patching:
        click.secho(f"\nPatching failed for {len(failed_patching)} slides", fg="yellow")
        click.secho("\n".join(failed_patching), fg="yellow")
    if failed_inference:
        click.secho(f"\nInference failed for {len(failed_inference)} slides", fg="yellow")
        click.secho("\n".join(failed_inference), fg="yellow")
    _barrier()
    if rank == 0:
        if geojson:
            click.echo("\nWriting inference results to GeoJSON files\n")
            csvs = sorted((Path(results_dir) / "model-outputs-csv").glob("*.csv"))
            write_geojsons(csvs=csvs, overlap=overlap, results_dir=Path(results_dir), output_dir="model-outputs-geojson")
        ts = datetime.now().astimezone().strftime("%Y%m%dT%H%M%S")
        out = Path(results_dir) / f"infer_metadata_{ts}.json"
        click.echo(f"\nSaving metadata about run to {out}\n")
        out.write_text(json.dumps(_run_metadata(model_obj), indent=2))
    _barrier()
    click.secho("\nWSInsight-infer tasks are all finished.\n", fg="green")


def _open_all(slide_paths, slide_mpp) -> Dict[str, object]:
    slides = {}
    for p in slide_paths:
        try:
            slides[str(p)] = open_slide(p, slide_mpp)
        except click.UsageError:
            raise
        except Exception as e:  # noqa: BLE001 - per-slide isolation like the reference's loops
            logging.getLogger(__name__).error(f"cannot open {p}: {e}")
    return slides


# ------------------------------------------------------------------------------------------------
# commands
# ------------------------------------------------------------------------------------------------
@click.group()
def cli():
    """wsinsight_b200: the WSInsight patch-classification hot path on B200."""
    logging.basicConfig(level=logging.INFO, format="%(levelname)s %(name)s: %(message)s")


def _prepare(kw):
    _unsupported(references_dir=kw.get("references_dir"))
    model_obj = _resolve_model(kw["model_name"], kw["config"], kw["model_path"], kw["qupath_detection_dir"],
                               kw["qupath_geojson_detection_dir"], kw["qupath_geojson_annotation_dir"])
    _init_distributed()
    slide_paths = enumerate_slide_paths(kw["wsi_dir"], kw.get("synthetic", ()))
    return model_obj, slide_paths


@cli.command()
@_common_options
@_patch_options
def patch(**kw):
    """Segment tissue and generate patch coordinates for a WSI directory."""
    _unsupported(cache_image_patches=kw["cache_image_patches"], histoqc_dir=kw["histoqc_dir"])
    model_obj, slide_paths = _prepare(kw)
    overlap = _overlap(model_obj.config, kw["patch_overlap_ratio"], kw["patch_size_um"], kw["patch_size_px"])
    rank, _ = _rank_world()
    slides = _open_all(slide_paths, kw["slide_mpp"]) if rank == 0 else {}
    do_patch(wsi_dir=kw["wsi_dir"], results_dir=kw["results_dir"], model_obj=model_obj, slide_paths=slide_paths,
             slides=slides, seg_thumbsize=kw["seg_thumbsize"], seg_median_filter_size=kw["seg_median_filter_size"],
             seg_binary_threshold=kw["seg_binary_threshold"], seg_otsu=kw["seg_otsu"],
             seg_closing_kernel_size=kw["seg_closing_kernel_size"], seg_closing_convention=kw["seg_closing_convention"],
             seg_min_object_size_um2=kw["seg_min_object_size_um2"], seg_min_hole_size_um2=kw["seg_min_hole_size_um2"],
             overlap=overlap)


@cli.command()
@_common_options
@_infer_options
@click.option("--patch-overlap-ratio", default=0.0, type=click.FloatRange(min=None, max=1, max_open=True))
@click.option("--patch-size-um", default=0.0, type=click.FloatRange(min=0.0))
@click.option("--patch-size-px", default=0, type=click.FloatRange(min=0))
def infer(**kw):
    """Classify the patches of RESULTS_DIR/patches and export CSV (+ GeoJSON)."""
    _unsupported(qupath=kw["qupath"], omecsv=kw["omecsv"], hplot=kw["hplot"], cme_cellular=kw["cme_cellular"],
                 cme_annotation=kw["cme_annotation"])
    model_obj, slide_paths = _prepare(kw)
    overlap = _overlap(model_obj.config, kw["patch_overlap_ratio"], kw["patch_size_um"], kw["patch_size_px"])
    slides = _open_all(slide_paths, kw["slide_mpp"])
    do_infer(wsi_dir=kw["wsi_dir"], results_dir=kw["results_dir"], model_obj=model_obj, slide_paths=slide_paths,
             slides=slides, batch_size=kw["batch_size"], num_workers=kw["num_workers"], geojson=kw["geojson"],
             overlap=overlap)


@cli.command()
@_common_options
@_infer_options
@_patch_options
def run(**kw):
    """Run both patch extraction and inference workflows for a slide directory."""
    _unsupported(cache_image_patches=kw["cache_image_patches"], histoqc_dir=kw["histoqc_dir"], qupath=kw["qupath"],
                 omecsv=kw["omecsv"], hplot=kw["hplot"], cme_cellular=kw["cme_cellular"],
                 cme_annotation=kw["cme_annotation"])
    model_obj, slide_paths = _prepare(kw)
    overlap = _overlap(model_obj.config, kw["patch_overlap_ratio"], kw["patch_size_um"], kw["patch_size_px"])
    slides = _open_all(slide_paths, kw["slide_mpp"])
    do_patch(wsi_dir=kw["wsi_dir"], results_dir=kw["results_dir"], model_obj=model_obj, slide_paths=slide_paths,
             slides=slides, seg_thumbsize=kw["seg_thumbsize"], seg_median_filter_size=kw["seg_median_filter_size"],
             seg_binary_threshold=kw["seg_binary_threshold"], seg_otsu=kw["seg_otsu"],
             seg_closing_kernel_size=kw["seg_closing_kernel_size"], seg_closing_convention=kw["seg_closing_convention"],
             seg_min_object_size_um2=kw["seg_min_object_size_um2"], seg_min_hole_size_um2=kw["seg_min_hole_size_um2"],
             overlap=overlap)
    do_infer(wsi_dir=kw["wsi_dir"], results_dir=kw["results_dir"], model_obj=model_obj, slide_paths=slide_paths,
             slides=slides, batch_size=kw["batch_size"], num_workers=kw["num_workers"], geojson=kw["geojson"],
             overlap=overlap)
    click.secho("\nWSInsight run completed.\n", fg="green")


def main(argv=None):
    try:
        cli.main(args=argv, prog_name="wsinsight_b200", standalone_mode=True)
    finally:
        try:
            import torch.distributed as dist

            if dist.is_available() and dist.is_initialized():
                dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            pass


if __name__ == "__main__":
    main()
