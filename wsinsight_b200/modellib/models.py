"""Model-zoo boundary: ``config.json`` + TorchScript file -> a B200 network.

Mirror of ``wsinsight/modellib/models.py:17-108``.  The reference loads the TorchScript module
lazily per device (``TSPerDevice``) and calls it through cuDNN.  Here the TorchScript file is
only a weight/topology container: its (frozen) graph is lowered once by
:mod:`wsinsight_b200.modellib.lowering` and executed by the tcgen05 kernels of
``libwsi_b200.so``.  ``wsinfer_zoo`` is not a dependency; :class:`ModelConfiguration` restates
the fields the reference reads (``tests/test_all.py:420-438``, ``cli/infer.py:835-848``).
"""

from __future__ import annotations

import dataclasses
import json
from pathlib import Path
from typing import Any, Dict, List, Optional

from ..errors import UnknownArchitectureError


@dataclasses.dataclass
class TransformConfigurationItem:
    name: str
    arguments: Optional[Dict[str, Any]] = None


@dataclasses.dataclass
class ModelConfiguration:
    """The wsinfer-zoo model configuration (spec_version 1.0)."""

    architecture: str
    num_classes: int
    class_names: List[str]
    patch_size_pixels: int
    spacing_um_px: float
    transform: List[TransformConfigurationItem]
    extras: Dict[str, Any] = dataclasses.field(default_factory=dict)

    @classmethod
    def from_dict(cls, d: Dict[str, Any]) -> "ModelConfiguration":
        required = ("architecture", "num_classes", "class_names", "patch_size_pixels", "spacing_um_px",
                    "transform")
        missing = [k for k in required if k not in d]
        if missing:
            raise ValueError(f"model configuration is missing keys: {missing}")
        if len(d["class_names"]) != int(d["num_classes"]):
            raise ValueError("length of class_names must equal num_classes")
        tf = [TransformConfigurationItem(t["name"], t.get("arguments")) for t in d["transform"]]
        extras = {k: v for k, v in d.items() if k not in required and k != "spec_version"}
        return cls(str(d["architecture"]), int(d["num_classes"]), list(d["class_names"]),
                   int(d["patch_size_pixels"]), float(d["spacing_um_px"]), tf, extras)

    @classmethod
    def from_json(cls, path) -> "ModelConfiguration":
        with open(path) as f:
            return cls.from_dict(json.load(f))


@dataclasses.dataclass
class Model:
    config: ModelConfiguration
    model_path: str


@dataclasses.dataclass
class LocalModelTorchScript(Model):
    """Typed stand-in for locally stored TorchScript bundles (models.py:17-21)."""


def get_registered_model(name: str):
    """models.py:24-36 resolves zoo names through ``wsinfer_zoo`` + HuggingFace; neither is
    available offline, and upstream's ``--model NAME`` branch is itself broken."""
    raise UnknownArchitectureError(
        f"registered-zoo lookup of {name!r} needs wsinfer_zoo and network access; "
        "use --config config.json --model-path torchscript_model.pt")


def get_pretrained_torch_module(model: Model, max_batch: int = 1024, device=None, input_size: int = None):
    """Reference name (models.py:93-108).  Returns a :class:`B200Network` (callable on the
    preprocess kernel's bf16 NHWC batches) instead of a cuDNN-backed TorchScript module."""
    from .engine import B200Network

    if not Path(model.model_path).exists():
        raise FileNotFoundError(model.model_path)
    if input_size is None:
        from .transforms import make_compose_from_transform_config

        input_size = make_compose_from_transform_config(model.config.transform).out_size(
            model.config.patch_size_pixels)
    return B200Network.from_torchscript(model.model_path, num_classes=model.config.num_classes,
                                        input_size=input_size, max_batch=max_batch, device=device)
