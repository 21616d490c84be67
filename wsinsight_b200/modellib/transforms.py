"""Model-config ``transform`` list -> parameters of the fused preprocess kernel.

Mirror of ``wsinsight/modellib/transforms.py:22-38``: the reference maps
``[{name, arguments}]`` onto ``torchvision.transforms.Compose([Resize, ToTensor, Normalize])``.
The B200 data path fuses those three into one kernel (csrc/preproc.cu), so the "compose" here
is a :class:`FusedTransform` record holding the resize target and the per-channel mean/std.
``Scale`` (min-max, transforms.py:8-18) is only used by cell models and is rejected.
"""

from __future__ import annotations

import dataclasses
from typing import Optional, Sequence

_SUPPORTED = ("Resize", "ToTensor", "Normalize")


@dataclasses.dataclass
class FusedTransform:
    size: Optional[int]              # Resize target (square patches: S x S); None = no resize
    mean: Optional[Sequence[float]]  # Normalize mean (3,), None = no Normalize
    std: Optional[Sequence[float]]

    def out_size(self, patch_size: int) -> int:
        return int(self.size) if self.size is not None else int(patch_size)


def make_compose_from_transform_config(list_of_transforms) -> FusedTransform:
    """Same name and input as transforms.py:29-38; returns the fused-kernel parameters."""
    size = mean = std = None
    seen = []
    for item in list_of_transforms:
        name = item.name if hasattr(item, "name") else item["name"]
        args = (item.arguments if hasattr(item, "arguments") else item.get("arguments")) or {}
        if name not in _SUPPORTED:
            raise KeyError(f"transform {name!r} is not supported by the fused B200 data path "
                           f"(supported: {', '.join(_SUPPORTED)})")
        seen.append(name)
        if name == "Resize":
            s = args["size"]
            if isinstance(s, (list, tuple)):
                if len(s) == 2 and s[0] != s[1]:
                    raise ValueError(f"non-square Resize {s} is not supported")
                s = s[0]
            size = int(s)
        elif name == "Normalize":
            mean = [float(v) for v in args["mean"]]
            std = [float(v) for v in args["std"]]
            if len(mean) != 3 or len(std) != 3:
                raise ValueError("Normalize needs 3-channel mean/std")
    if "ToTensor" not in seen:
        raise ValueError("the transform list must contain ToTensor")
    order = [n for n in seen]
    if order != sorted(order, key=_SUPPORTED.index):
        raise ValueError(f"transform order {order} differs from Resize -> ToTensor -> Normalize")
    return FusedTransform(size, mean, std)
