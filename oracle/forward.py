"""Oracle for the model forward + head (TEST INFRASTRUCTURE ONLY).

Follows the reference classification loop wsinsight/modellib/run_inference.py:537-560:
``logits = model(x)`` in fp32, then ``softmax(dim=1)`` if K>1 else
``sigmoid(logits.squeeze(1))`` on the CPU.

Model topologies: ResNet34 = torchvision ``resnet34``
with ``fc -> K`` (state-dict keys in the reference's scripts/README.md:31);
zoo files are produced by ``torch.jit.script`` + ``torch.jit.freeze``
(scripts/convert_to_torchscript.py:38-50).  The real zoo weights are not reachable
offline, so parity is on identical random-init weights ("parity unpinned" vs the
golden CSV probabilities).
"""

from __future__ import annotations

import json
from pathlib import Path

import torch
import torchvision


def make_resnet34(num_classes: int = 2, seed: int = 0) -> torch.nn.Module:
    """torchvision resnet34, default init, BN running stats randomised so folding matters."""
    g = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    m = torchvision.models.resnet34(num_classes=num_classes)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.running_mean.copy_(torch.randn(mod.running_mean.shape, generator=g) * 0.1)
            mod.running_var.copy_(torch.rand(mod.running_var.shape, generator=g) + 0.5)
            with torch.no_grad():
                mod.weight.copy_(torch.rand(mod.weight.shape, generator=g) * 0.5 + 0.75)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
    return m.eval()


def make_resnet34_calibrated(num_classes: int = 2, seed: int = 0, fc_gain: float = 4.0) -> torch.nn.Module:
    """ResNet34 with trained-net-like statistics: random affine BN parameters, running stats
    calibrated on random normalised-image-like batches (every layer O(1)), classifier scaled so the
    logits are O(1) and the probabilities are NOT saturated.  This is the hard case for bf16."""
    g = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    m = torchvision.models.resnet34(num_classes=num_classes)
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            with torch.no_grad():
                mod.weight.copy_(torch.rand(mod.weight.shape, generator=g) * 0.5 + 0.75)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.1)
            mod.momentum = None
            mod.reset_running_stats()
    m.train()
    with torch.no_grad():
        for _ in range(2):
            m(torch.randn((16, 3, 224, 224), generator=g) * 1.5)
        m.fc.weight.mul_(fc_gain)
    return m.eval()


BREAST_RESNET34_CONFIG = {
    "spec_version": "1.0",
    "architecture": "resnet34",
    "num_classes": 2,
    "class_names": ["Other", "Tumor"],
    "patch_size_pixels": 350,
    "spacing_um_px": 0.25,
    "transform": [
        {"name": "Resize", "arguments": {"size": 224}},
        {"name": "ToTensor"},
        {"name": "Normalize", "arguments": {"mean": [0.7238, 0.5716, 0.6779],
                                            "std": [0.112, 0.1459, 0.1089]}},
    ],
}


def export_torchscript(model: torch.nn.Module, path, frozen: bool = True) -> None:
    """scripts/convert_to_torchscript.py:38-50: jit.script (+ freeze) + save."""
    ts = torch.jit.script(model.eval())
    if frozen:
        ts = torch.jit.freeze(ts)
    torch.jit.save(ts, str(path))


def write_model_bundle(dirpath, num_classes=2, seed=0, frozen=True, config=None):
    """Write config.json + torchscript_model.pt the way a zoo entry ships them."""
    d = Path(dirpath)
    d.mkdir(parents=True, exist_ok=True)
    m = make_resnet34(num_classes, seed)
    export_torchscript(m, d / "torchscript_model.pt", frozen)
    cfg = dict(config or BREAST_RESNET34_CONFIG)
    cfg["num_classes"] = num_classes
    if num_classes != len(cfg["class_names"]):
        cfg["class_names"] = [f"c{i}" for i in range(num_classes)]
    (d / "config.json").write_text(json.dumps(cfg))
    return m, d / "config.json", d / "torchscript_model.pt"


@torch.no_grad()
def head(logits: torch.Tensor) -> torch.Tensor:
    """run_inference.py:548-551."""
    if logits.dim() > 1 and logits.shape[1] > 1:
        return torch.nn.functional.softmax(logits, dim=1)
    return torch.sigmoid(logits.squeeze(1))


@torch.no_grad()
def forward_probs(model: torch.nn.Module, x_nchw_f32: torch.Tensor) -> torch.Tensor:
    """fp32 forward + head on the CPU (the reference path under WSINFER_FORCE_CPU)."""
    return head(model(x_nchw_f32.float()))
