"""Oracle: InceptionV4 without BatchNorm (TEST INFRASTRUCTURE ONLY).

The reference ships no model source; the only in-repo description of the TIL lymphocyte model
(``pancancer-lymphocytes-inceptionv4.tcga``) is its weight converter
``scripts/convert_tf_to_pytorch_til_inceptionv4.py:33-206``: Cadene's InceptionV4 module paths
(``features.0 .. features.21``, ``last_linear``) where every ``BasicConv2d`` is a biased
convolution followed by ReLU and there is no BatchNorm.  This file restates that layout
(149 convolutions, 12.25 GMAC at 299^2) so that the B200 lowering and
kernels can be checked against a plain fp32 torch forward on identical random-init weights.
The real zoo weights are not reachable offline: "parity unpinned" vs the golden CSV probabilities.
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


class BasicConv2d(nn.Module):
    """conv (with bias) + ReLU -- the no-BatchNorm variant the TIL converter fills (:35-204)."""

    def __init__(self, cin, cout, kernel_size, stride=1, padding=0):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, kernel_size=kernel_size, stride=stride, padding=padding, bias=True)

    def forward(self, x):
        return F.relu(self.conv(x))


class Mixed_3a(nn.Module):
    def __init__(self):
        super().__init__()
        self.maxpool = nn.MaxPool2d(3, stride=2)
        self.conv = BasicConv2d(64, 96, 3, stride=2)

    def forward(self, x):
        return torch.cat([self.maxpool(x), self.conv(x)], 1)


class Mixed_4a(nn.Module):
    def __init__(self):
        super().__init__()
        self.branch0 = nn.Sequential(BasicConv2d(160, 64, 1), BasicConv2d(64, 96, 3))
        self.branch1 = nn.Sequential(BasicConv2d(160, 64, 1), BasicConv2d(64, 64, (1, 7), padding=(0, 3)),
                                     BasicConv2d(64, 64, (7, 1), padding=(3, 0)), BasicConv2d(64, 96, 3))

    def forward(self, x):
        return torch.cat([self.branch0(x), self.branch1(x)], 1)


class Mixed_5a(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = BasicConv2d(192, 192, 3, stride=2)
        self.maxpool = nn.MaxPool2d(3, stride=2)

    def forward(self, x):
        return torch.cat([self.conv(x), self.maxpool(x)], 1)


class Inception_A(nn.Module):
    def __init__(self):
        super().__init__()
        self.branch0 = BasicConv2d(384, 96, 1)
        self.branch1 = nn.Sequential(BasicConv2d(384, 64, 1), BasicConv2d(64, 96, 3, padding=1))
        self.branch2 = nn.Sequential(BasicConv2d(384, 64, 1), BasicConv2d(64, 96, 3, padding=1),
                                     BasicConv2d(96, 96, 3, padding=1))
        self.branch3 = nn.Sequential(nn.AvgPool2d(3, stride=1, padding=1, count_include_pad=False),
                                     BasicConv2d(384, 96, 1))

    def forward(self, x):
        return torch.cat([self.branch0(x), self.branch1(x), self.branch2(x), self.branch3(x)], 1)


class Reduction_A(nn.Module):
    def __init__(self):
        super().__init__()
        self.branch0 = BasicConv2d(384, 384, 3, stride=2)
        self.branch1 = nn.Sequential(BasicConv2d(384, 192, 1), BasicConv2d(192, 224, 3, padding=1),
                                     BasicConv2d(224, 256, 3, stride=2))
        self.branch2 = nn.MaxPool2d(3, stride=2)

    def forward(self, x):
        return torch.cat([self.branch0(x), self.branch1(x), self.branch2(x)], 1)


class Inception_B(nn.Module):
    def __init__(self):
        super().__init__()
        self.branch0 = BasicConv2d(1024, 384, 1)
        self.branch1 = nn.Sequential(BasicConv2d(1024, 192, 1), BasicConv2d(192, 224, (1, 7), padding=(0, 3)),
                                     BasicConv2d(224, 256, (7, 1), padding=(3, 0)))
        self.branch2 = nn.Sequential(BasicConv2d(1024, 192, 1), BasicConv2d(192, 192, (7, 1), padding=(3, 0)),
                                     BasicConv2d(192, 224, (1, 7), padding=(0, 3)),
                                     BasicConv2d(224, 224, (7, 1), padding=(3, 0)),
                                     BasicConv2d(224, 256, (1, 7), padding=(0, 3)))
        self.branch3 = nn.Sequential(nn.AvgPool2d(3, stride=1, padding=1, count_include_pad=False),
                                     BasicConv2d(1024, 128, 1))

    def forward(self, x):
        return torch.cat([self.branch0(x), self.branch1(x), self.branch2(x), self.branch3(x)], 1)


class Reduction_B(nn.Module):
    def __init__(self):
        super().__init__()
        self.branch0 = nn.Sequential(BasicConv2d(1024, 192, 1), BasicConv2d(192, 192, 3, stride=2))
        self.branch1 = nn.Sequential(BasicConv2d(1024, 256, 1), BasicConv2d(256, 256, (1, 7), padding=(0, 3)),
                                     BasicConv2d(256, 320, (7, 1), padding=(3, 0)), BasicConv2d(320, 320, 3, stride=2))
        self.branch2 = nn.MaxPool2d(3, stride=2)

    def forward(self, x):
        return torch.cat([self.branch0(x), self.branch1(x), self.branch2(x)], 1)


class Inception_C(nn.Module):
    def __init__(self):
        super().__init__()
        self.branch0 = BasicConv2d(1536, 256, 1)
        self.branch1_0 = BasicConv2d(1536, 384, 1)
        self.branch1_1a = BasicConv2d(384, 256, (1, 3), padding=(0, 1))
        self.branch1_1b = BasicConv2d(384, 256, (3, 1), padding=(1, 0))
        self.branch2_0 = BasicConv2d(1536, 384, 1)
        self.branch2_1 = BasicConv2d(384, 448, (3, 1), padding=(1, 0))
        self.branch2_2 = BasicConv2d(448, 512, (1, 3), padding=(0, 1))
        self.branch2_3a = BasicConv2d(512, 256, (1, 3), padding=(0, 1))
        self.branch2_3b = BasicConv2d(512, 256, (3, 1), padding=(1, 0))
        self.branch3 = nn.Sequential(nn.AvgPool2d(3, stride=1, padding=1, count_include_pad=False),
                                     BasicConv2d(1536, 256, 1))

    def forward(self, x):
        x0 = self.branch0(x)
        x1 = self.branch1_0(x)
        x1a = self.branch1_1a(x1)
        x1b = self.branch1_1b(x1)
        x2 = self.branch2_2(self.branch2_1(self.branch2_0(x)))
        x2a = self.branch2_3a(x2)
        x2b = self.branch2_3b(x2)
        return torch.cat([x0, x1a, x1b, x2a, x2b, self.branch3(x)], 1)


class InceptionV4NoBN(nn.Module):
    """features.0-21 + 8x8 average pool + last_linear (converter :33-206)."""

    def __init__(self, num_classes: int = 2):
        super().__init__()
        self.features = nn.Sequential(
            BasicConv2d(3, 32, 3, stride=2), BasicConv2d(32, 32, 3), BasicConv2d(32, 64, 3, padding=1),
            Mixed_3a(), Mixed_4a(), Mixed_5a(),
            Inception_A(), Inception_A(), Inception_A(), Inception_A(), Reduction_A(),
            Inception_B(), Inception_B(), Inception_B(), Inception_B(), Inception_B(), Inception_B(), Inception_B(),
            Reduction_B(), Inception_C(), Inception_C(), Inception_C())
        self.last_linear = nn.Linear(1536, num_classes)

    def forward(self, x):
        x = self.features(x)
        x = F.adaptive_avg_pool2d(x, (1, 1))
        x = torch.flatten(x, 1)
        return self.last_linear(x)


TIL_INCEPTIONV4_CONFIG = {
    "spec_version": "1.0",
    "architecture": "inception_v4",
    "num_classes": 2,
    "class_names": ["Other", "Lymphocytes"],
    "patch_size_pixels": 100,
    "spacing_um_px": 0.5,
    "transform": [
        {"name": "Resize", "arguments": {"size": 299}},
        {"name": "ToTensor"},
        {"name": "Normalize", "arguments": {"mean": [0.5, 0.5, 0.5], "std": [0.5, 0.5, 0.5]}},
    ],
}


def make_inceptionv4(num_classes: int = 2, seed: int = 0, fc_gain: float = 1.0) -> nn.Module:
    """Random-init model with variance-preserving weights (He normal; there is no BatchNorm to rescale),
    small random biases, so that every layer stays O(1) and the logits depend on the input."""
    torch.manual_seed(seed)
    m = InceptionV4NoBN(num_classes)
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, nn.Conv2d):
                fan_in = mod.in_channels * mod.kernel_size[0] * mod.kernel_size[1]
                mod.weight.copy_(torch.randn(mod.weight.shape, generator=g) * (2.0 / fan_in) ** 0.5)
                mod.bias.copy_(torch.randn(mod.bias.shape, generator=g) * 0.05)
        m.last_linear.weight.mul_(fc_gain)
    return m.eval()
