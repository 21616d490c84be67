"""TEST INFRASTRUCTURE ONLY -- the reference's tile-stitcher batch step on the CPU.

Executes the very torch calls of ``wsinsight/modellib/tilefuse.py:240-278`` (softmax, ``F.interpolate`` bilinear
``align_corners=False``, ``* alpha``, renormalisation, clipped paste at ``coords + halo``) on CPU tensors; these
ARE the reference's operations, so the pin is the library itself.  fp32: compared within 2e-6 absolute.
"""

from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F


def accumulate(np_logits, hv, tp_logits, coords, S, halo, alpha, W, H, np_map=None, hv_map=None, tp_map=None):
    K = tp_logits.shape[1]
    np_map = np.zeros((H, W), np.float32) if np_map is None else np_map
    hv_map = np.zeros((H, W, 2), np.float32) if hv_map is None else hv_map
    tp_map = np.zeros((H, W, K), np.float32) if tp_map is None else tp_map
    np_prob = torch.softmax(np_logits, dim=1)[:, 1:2]                                   # :240
    tp_prob = torch.softmax(tp_logits, dim=1)                                           # :241
    np_res = F.interpolate(np_prob, size=(S, S), mode="bilinear", align_corners=False)  # :244
    hv_res = F.interpolate(hv, size=(S, S), mode="bilinear", align_corners=False) * alpha
    tp_res = F.interpolate(tp_prob, size=(S, S), mode="bilinear", align_corners=False)
    tp_res = tp_res / (tp_res.sum(dim=1, keepdim=True) + 1e-8)                          # :252
    a = np_res.squeeze(1).numpy()
    b = hv_res.permute(0, 2, 3, 1).numpy()
    c = tp_res.permute(0, 2, 3, 1).numpy()
    xy = np.asarray(coords)[:, :2].astype(np.int32) + halo                              # :259-260
    for i in range(len(xy)):                                                             # :263-278
        x0, y0 = int(xy[i, 0]), int(xy[i, 1])
        cx0, cy0, cx1, cy1 = max(0, x0), max(0, y0), min(W, x0 + S), min(H, y0 + S)
        if cx1 <= cx0 or cy1 <= cy0:
            continue
        tx0, ty0 = cx0 - x0, cy0 - y0
        np_map[cy0:cy1, cx0:cx1] = a[i, ty0:ty0 + cy1 - cy0, tx0:tx0 + cx1 - cx0]
        hv_map[cy0:cy1, cx0:cx1] = b[i, ty0:ty0 + cy1 - cy0, tx0:tx0 + cx1 - cx0]
        tp_map[cy0:cy1, cx0:cx1] = c[i, ty0:ty0 + cy1 - cy0, tx0:tx0 + cx1 - cx0]
    return np_map, hv_map, tp_map
