"""B200 network: lowered IR -> ``wsi_net_*`` handle -> probabilities.

Replaces the TorchScript forward + CPU softmax of the reference classification loop
(``wsinsight/modellib/run_inference.py:537-560``).  The object is callable like the module the
reference gets from ``get_pretrained_torch_module`` (``models.py:93-108``) but takes the bf16
NHWC batches written by the fused preprocess kernel and returns fp32 probabilities
(softmax for K > 1, sigmoid for K == 1) computed on the device.
"""

from __future__ import annotations

import ctypes as C
from typing import List

import numpy as np

from .. import _cabi
from ..errors import UnknownArchitectureError
from .lowering import IRNode, lower_torchscript


def _fptr(t):
    """ctypes float* of a contiguous fp32 CPU tensor (or None)."""
    if t is None:
        return None
    return C.cast(t.data_ptr(), _cabi.c_float_p)


def _topo_order(nodes: List[IRNode]) -> List[int]:
    order, state = [], [0] * len(nodes)

    def visit(i):
        stack = [(i, 0)]
        while stack:
            j, phase = stack.pop()
            if phase == 0:
                if state[j]:
                    continue
                state[j] = 1
                stack.append((j, 1))
                deps = list(nodes[j].inputs) + ([nodes[j].residual] if nodes[j].residual is not None else [])
                for d in reversed(deps):
                    if not state[d]:
                        stack.append((d, 0))
            else:
                order.append(j)

    for i in range(len(nodes)):
        visit(i)
    return order


class B200Network:
    """A classification CNN resident on one B200, executed only by libwsi_b200 kernels."""

    def __init__(self, nodes: List[IRNode], max_batch: int, device=None):
        torch = _cabi.require_cuda()
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.max_batch = int(max_batch)
        self.nodes = nodes
        self._h = C.c_void_p()
        lib = _cabi.load()
        with torch.cuda.device(self.device):
            _cabi.call("wsi_net_create", self.max_batch, C.byref(self._h))
            try:
                self._build(nodes)
                _cabi.call("wsi_net_finalize", self._h)
            except Exception:
                lib.wsi_net_destroy(self._h)
                self._h = C.c_void_p()
                raise
        info = (C.c_int * 4)()
        _cabi.call("wsi_net_input_layout", self._h, info)
        self.input_bytes = int(lib.wsi_net_input_bytes(self._h))
        if int(info[0]) == 1:
            # row-parity planes of pixel pairs (csrc/conv_stem.cuh): bf16 [2][max_batch][P][P][8]
            self.input_layout.update(kind="parity", P=int(info[1]),
                                     plane_stride_px=(int(info[3]) << 32) | (int(info[2]) & 0xFFFFFFFF))
        else:
            self.input_layout.update(kind="nhwc4", plane_stride_px=0)
        self.input_layout["buffer_elems"] = self.input_bytes // 2
        self.num_launches = int(lib.wsi_net_num_launches(self._h))
        self.num_ops = int(lib.wsi_net_num_ops(self._h))
        self.flops_per_image = float(lib.wsi_net_flops_per_image(self._h))

    # -- construction ---------------------------------------------------------
    @classmethod
    def from_torchscript(cls, path_or_module, num_classes=None, input_size: int = 224, max_batch: int = 1024,
                         device=None):
        nodes = lower_torchscript(path_or_module, input_size)
        net = cls(nodes, max_batch, device)
        if num_classes is not None and net.num_classes != int(num_classes):
            raise UnknownArchitectureError(
                f"model emits {net.num_classes} classes but the configuration says {num_classes}")
        return net

    def _add_tensor(self, h, w, c, external=False) -> int:
        tid = C.c_int()
        _cabi.call("wsi_net_add_tensor", self._h, int(h), int(w), int(c), 1 if external else 0, C.byref(tid))
        return tid.value

    def _build(self, nodes):
        inp = nodes[0]
        self.input_size = int(inp.shape[1])
        first_convs = [n for n in nodes if n.kind == "conv" and n.inputs == [0]]
        if len(first_convs) != 1 or nodes[0].n_uses != 1:
            raise UnknownArchitectureError("the network input must feed exactly one convolution")
        stem = first_convs[0]
        kh, kw = int(stem.weight.shape[2]), int(stem.weight.shape[3])
        S = self.input_size
        stem7 = (kh == 7 and kw == 7 and stem.stride == 2 and tuple(stem.pad) == (3, 3) and stem.residual is None
                 and S % 2 == 0)
        if stem.residual is not None:
            raise UnknownArchitectureError("the first convolution cannot take a residual")
        import os

        # InceptionV4 conv2d_1a: 3x3 / stride 2 / no padding, 32 (or 64) filters -> the same parity-plane kernel
        stem3 = (kh == 3 and kw == 3 and stem.stride == 2 and tuple(stem.pad) == (0, 0)
                 and int(stem.weight.shape[0]) in (32, 64) and os.environ.get("WSI_B200_NO_FLAT") != "1")
        tensor_of, c_off = {}, {}
        if stem3:
            se = S + (S & 1)       # stored side rounded up to even; the extra row / column stays zero and is never used
            self.input_layout = dict(out_H=se, out_W=se, pad_top=0, pad_left=0, channels=4)
            tensor_of[0] = self._add_tensor(se, se, 4, external=True)
        elif stem7:
            # ResNet family: the preprocess kernel writes the padded image, patch at (3,3), 4 channels
            # (as row-parity planes when the 64-filter stem kernel is used, see csrc/conv_stem.cuh)
            self.input_layout = dict(out_H=S + 6, out_W=S + 6, pad_top=3, pad_left=3, channels=4)
            tensor_of[0] = self._add_tensor(S + 6, S + 6, 4, external=True)
        else:
            # any other first convolution (InceptionV4: 3x3/2, no padding) runs on the generic per-tap kernel
            # over an 8-channel NHWC image (channels 3..7 zero, weights zero-padded to match)
            self.input_layout = dict(out_H=S, out_W=S, pad_top=0, pad_left=0, channels=8)
            tensor_of[0] = self._add_tensor(S, S, 8, external=True)
        c_off[0] = 0
        for i, n in enumerate(nodes):
            if n.kind == "cat":
                c, h, w = n.shape
                tid = self._add_tensor(h, w, c)
                tensor_of[i], c_off[i] = tid, 0
                off = 0
                for p in n.inputs:
                    if nodes[p].kind not in ("conv", "pool") or nodes[p].n_uses != 1 or p in tensor_of:
                        raise UnknownArchitectureError("cat input must be a single-use conv/pool output")
                    if off % 8 or nodes[p].shape[0] % 8:
                        raise UnknownArchitectureError("cat channel offsets must be multiples of 8")
                    tensor_of[p], c_off[p] = tid, off
                    off += nodes[p].shape[0]
        # Pool branch of an Inception block: avg_pool2d(3, 1, 1) -> 1x1 conv.  Both are linear and the pool divisor
        # depends on the position only, so conv1x1(avgpool(x)) == avgpool(conv1x1(x)): run the conv FIRST (no bias, no
        # ReLU; it reads x like its sibling branches) and pool its 4-8x narrower output with bias + ReLU fused into the
        # pooling kernel (wsi_net_add_pool_bias).  The wide block input is then no longer read and re-written by a
        # pooling pass.  Pool node i keeps its index but now holds conv1x1_nobias(x) (self.commuted_pools).
        commute = {}
        if os.environ.get("WSI_B200_NO_POOL_COMMUTE") != "1":
            for j, m in enumerate(nodes):
                if (m.kind == "conv" and m is not stem and m.residual is None and tuple(m.weight.shape[2:]) == (1, 1)
                        and int(m.stride) == 1 and tuple(m.pad) == (0, 0) and int(m.weight.shape[0]) % 32 == 0):
                    i = m.inputs[0]
                    p = nodes[i]
                    if (p.kind == "pool" and not p.pool_max and int(p.k) == 3 and int(p.pstride) == 1
                            and int(p.ppad) == 1 and p.n_uses == 1 and i not in tensor_of
                            and int(m.weight.shape[0]) < int(p.shape[0])):
                        commute[i] = j
        self.commuted_pools = sorted(commute)
        conv_of_pool = dict(commute)
        pool_of_conv = {j: i for i, j in commute.items()}
        # Sibling 1x1 convolutions (same input, stride 1, no padding, same activation -- the branch heads of an
        # Inception block) whose outputs feed only convolutions: ONE launch with the filters stacked writes one wide
        # tensor and the consumers read channel slices of it, so the block input is read from HBM once per group
        # instead of once per branch (these 1x1 convs are HBM-bound: 384 -> 64 at 35x35 moves 1.1 GB for 0.05 GFLOP/img).
        merged_into, group_of = {}, {}
        users = {}
        for j, m in enumerate(nodes):
            for p in m.inputs:
                users.setdefault(p, []).append((j, "in"))
            if getattr(m, "residual", None) is not None:
                users.setdefault(m.residual, []).append((j, "res"))
        if os.environ.get("WSI_B200_NO_SIBLING_MERGE") != "1":
            sibs = {}
            for i, n in enumerate(nodes):
                if (n.kind == "conv" and i not in tensor_of and n is not stem and n.residual is None
                        and tuple(n.weight.shape[2:]) == (1, 1) and int(n.stride) == 1 and tuple(n.pad) == (0, 0)
                        and int(n.weight.shape[0]) % 8 == 0
                        and all(nodes[j].kind == "conv" and how == "in" for j, how in users.get(i, []))):
                    sibs.setdefault((n.inputs[0], bool(n.relu)), []).append(i)
            # the bias-free, ReLU-free 1x1 conv of a commuted pool branch reads the same input: it rides in the same
            # launch as the LAST slice (ReLU only on the channels before it: wsi_net_set_conv_relu_cols)
            for pi in sorted(conv_of_pool):
                key = (nodes[pi].inputs[0], True)
                if key in sibs:
                    sibs[key].append(pi)

            def mcout(m):
                return int(nodes[conv_of_pool[m]].weight.shape[0]) if nodes[m].kind == "pool" else int(nodes[m].weight.shape[0])

            for (x, _), members in sibs.items():
                relu_c = sum(mcout(m) for m in members if nodes[m].kind == "conv")
                if any(nodes[m].kind == "pool" for m in members) and relu_c % 64:
                    members = [m for m in members if nodes[m].kind == "conv"]
                if len(members) < 2 or sum(mcout(m) for m in members) % 32:
                    continue
                c, h, w = nodes[members[0]].shape
                tid = self._add_tensor(h, w, sum(mcout(m) for m in members))
                off = 0
                for i in members:
                    tensor_of[i], c_off[i] = tid, off
                    off += mcout(i)
                    merged_into[i] = members[0]
                group_of[members[0]] = members
        self.merged_groups = list(group_of.values())
        # Downsample branch of a torchvision BasicBlock: conv1x1(stride s, no padding) reads the pixels the centre tap of
        # the block's first conv3x3(stride s, padding 1) reads.  Its filters are stacked under the 3x3 ones (zero outside
        # the centre tap; the kernel skips the other taps for those channels: wsi_net_set_conv_center_cols), one launch
        # writes one tensor [main | downsample], and the block's second conv takes its input and its residual as the
        # two channel slices of that tensor.  Same operands and rounding points as two launches.
        ds_main, ds_of = {}, {}                       # downsample node -> main conv node, and the inverse
        if os.environ.get("WSI_B200_NO_DS_MERGE") != "1":
            for i, n in enumerate(nodes):
                if not (n.kind == "conv" and i not in tensor_of and n is not stem and n.residual is None and not n.relu
                        and tuple(n.weight.shape[2:]) == (1, 1) and int(n.stride) > 1 and tuple(n.pad) == (0, 0)
                        and int(n.weight.shape[0]) % 64 == 0 and users.get(i)
                        and all(how == "res" and nodes[j].kind == "conv" for j, how in users[i])):
                    continue
                for j, m in enumerate(nodes):
                    if (j != i and m.kind == "conv" and m.inputs == n.inputs and j not in tensor_of and j not in ds_of
                            and m is not stem and m.residual is None and m.relu
                            and tuple(m.weight.shape[2:]) == (3, 3) and int(m.stride) == int(n.stride)
                            and tuple(m.pad) == (1, 1) and tuple(m.shape[1:]) == tuple(n.shape[1:])
                            and int(m.weight.shape[0]) % 64 == 0 and users.get(j)
                            and all(how == "in" and nodes[k].kind == "conv" for k, how in users[j])):
                        c, h, w = m.shape
                        tid = self._add_tensor(h, w, int(m.weight.shape[0]) + int(n.weight.shape[0]))
                        tensor_of[j], c_off[j] = tid, 0
                        tensor_of[i], c_off[i] = tid, int(m.weight.shape[0])
                        ds_main[i], ds_of[j] = j, i
                        break
        self.downsample_merges = sorted(ds_of.items())
        for i, n in enumerate(nodes):
            if n.kind in ("conv", "pool") and i not in tensor_of:
                c, h, w = n.shape
                if i in conv_of_pool:
                    c = int(nodes[conv_of_pool[i]].weight.shape[0])
                tensor_of[i], c_off[i] = self._add_tensor(h, w, c), 0
        self.num_classes = None
        self._keep = []
        emitted_groups = set()
        for i in _topo_order(nodes):
            n = nodes[i]
            if n.kind in ("input", "cat"):
                continue
            x = n.inputs[0]
            in_off = 0
            if x in merged_into or x in ds_of:
                in_off = c_off[x]                  # slice of a merged sibling tensor (conv consumers only)
            elif c_off.get(x, 0) != 0 or (nodes[x].kind in ("conv", "pool") and
                                          nodes[x].shape[0] != self._tensor_channels(nodes, tensor_of, x)):
                raise UnknownArchitectureError("reading a channel slice of a concat buffer is not supported")
            if n.kind == "conv" and i in pool_of_conv:
                b = (n.bias if n.bias is not None else __import__("torch").zeros(int(n.weight.shape[0]))).contiguous()
                self._keep.append(b)
                pn = nodes[pool_of_conv[i]]
                _cabi.call("wsi_net_add_pool_bias", self._h, tensor_of[x], int(in_off), int(n.weight.shape[0]),
                           tensor_of[i], c_off[i], int(pn.k), int(pn.pstride), int(pn.ppad),
                           1 if pn.count_include_pad else 0, _fptr(b), 1 if n.relu else 0)
                continue
            if i in merged_into:
                # one launch for the whole sibling group, emitted with its first visited member
                import torch

                if merged_into[i] in emitted_groups:
                    continue
                emitted_groups.add(merged_into[i])
                ws, bs, relu_c, has_pool = [], [], 0, False
                for m in group_of[merged_into[i]]:
                    mn = nodes[conv_of_pool[m]] if nodes[m].kind == "pool" else nodes[m]
                    ws.append(mn.weight)
                    if nodes[m].kind == "pool":
                        has_pool = True
                        bs.append(torch.zeros(int(mn.weight.shape[0])))
                    else:
                        relu_c += int(mn.weight.shape[0])
                        bs.append(mn.bias if mn.bias is not None else torch.zeros(int(mn.weight.shape[0])))
                w, b = torch.cat(ws, 0).contiguous(), torch.cat(bs, 0).contiguous()
                self._keep += [w, b]
                first = nodes[group_of[merged_into[i]][0]]
                _cabi.call("wsi_net_add_conv_slice", self._h, tensor_of[x], int(in_off), tensor_of[i], 0, _fptr(w),
                           _fptr(b), int(w.shape[1]), int(w.shape[0]), 1, 1, 1, 0, 0, 1 if first.relu else 0, -1)
                if has_pool:
                    _cabi.call("wsi_net_set_conv_relu_cols", self._h, relu_c)
                continue
            if i in ds_main or i in ds_of:
                # block conv1 + downsample: one launch, emitted with whichever node is visited first
                import torch

                j = i if i in ds_of else ds_main[i]
                if j in emitted_groups:
                    continue
                emitted_groups.add(j)
                mn, dn = nodes[j], nodes[ds_of[j]]
                cm, cd, cin = int(mn.weight.shape[0]), int(dn.weight.shape[0]), int(mn.weight.shape[1])
                wd = torch.zeros((cd, cin, 3, 3), dtype=torch.float32)
                wd[:, :, 1, 1] = dn.weight[:, :, 0, 0]
                w = torch.cat([mn.weight, wd], 0).contiguous()
                b = torch.cat([mn.bias if mn.bias is not None else torch.zeros(cm),
                               dn.bias if dn.bias is not None else torch.zeros(cd)], 0).contiguous()
                self._keep += [w, b]
                _cabi.call("wsi_net_add_conv_slice", self._h, tensor_of[x], int(in_off), tensor_of[j], 0, _fptr(w),
                           _fptr(b), cin, cm + cd, 3, 3, int(mn.stride), 1, 1, 1, -1)
                _cabi.call("wsi_net_set_conv_relu_cols", self._h, cm)
                _cabi.call("wsi_net_set_conv_center_cols", self._h, cm)
                continue
            if n.kind == "pool" and i in conv_of_pool:
                cn = nodes[conv_of_pool[i]]
                w = cn.weight.contiguous()
                self._keep.append(w)
                _cabi.call("wsi_net_add_conv_slice", self._h, tensor_of[x], int(in_off), tensor_of[i], 0, _fptr(w), None,
                           int(w.shape[1]), int(w.shape[0]), 1, 1, 1, 0, 0, 0, -1)
                continue
            if n.kind == "conv":
                cout, cin = int(n.weight.shape[0]), int(n.weight.shape[1])
                w = n.weight.contiguous()
                b = n.bias.contiguous() if n.bias is not None else None
                out_off = c_off[i]
                self._keep += [w, b]
                if n is stem and stem3:
                    _cabi.call("wsi_net_add_stem_parity", self._h, tensor_of[x], tensor_of[i], _fptr(w), _fptr(b), cout,
                               3, 0, 1 if n.relu else 0)
                elif n is stem and not stem7:
                    import torch

                    w8 = torch.zeros((cout, 8, int(w.shape[2]), int(w.shape[3])), dtype=torch.float32)
                    w8[:, :cin] = w
                    self._keep.append(w8)
                    _cabi.call("wsi_net_add_conv", self._h, tensor_of[x], tensor_of[i], c_off[i], _fptr(w8), _fptr(b),
                               8, cout, int(w.shape[2]), int(w.shape[3]), int(n.stride), int(n.pad[0]),
                               int(n.pad[1]), 1 if n.relu else 0, -1)
                elif n is stem:
                    _cabi.call("wsi_net_add_stem7x7", self._h, tensor_of[x], tensor_of[i], _fptr(w), _fptr(b), cout,
                               1 if n.relu else 0)
                else:
                    res = tensor_of[n.residual] if n.residual is not None else -1
                    res_slice = n.residual is not None and n.residual in ds_main
                    if n.residual is not None and c_off.get(n.residual, 0) != 0 and not res_slice:
                        raise UnknownArchitectureError("residual read from a concat slice")
                    _cabi.call("wsi_net_add_conv_slice", self._h, tensor_of[x], int(in_off), tensor_of[i], out_off,
                               _fptr(w), _fptr(b), cin, cout, int(w.shape[2]), int(w.shape[3]), int(n.stride),
                               int(n.pad[0]), int(n.pad[1]), 1 if n.relu else 0, -1 if res_slice else res)
                    if res_slice:
                        _cabi.call("wsi_net_set_conv_residual_slice", self._h, res, int(c_off[n.residual]))
            elif n.kind == "pool":
                _cabi.call("wsi_net_add_pool", self._h, tensor_of[x], tensor_of[i], c_off[i], 1 if n.pool_max else 0,
                           int(n.k), int(n.pstride), int(n.ppad), 1 if n.count_include_pad else 0)
            elif n.kind == "head":
                w = n.fc_w.contiguous()
                b = n.fc_b.contiguous() if n.fc_b is not None else None
                self._keep += [w, b]
                self.num_classes = int(w.shape[0])
                _cabi.call("wsi_net_add_head", self._h, tensor_of[x], _fptr(w), _fptr(b), self.num_classes)
        if self.num_classes is None:
            raise UnknownArchitectureError("no classifier head found")
        self._tensor_of = tensor_of
        self._c_off = c_off

    @staticmethod
    def _tensor_channels(nodes, tensor_of, x):
        tid = tensor_of[x]
        for i, n in enumerate(nodes):
            if tensor_of.get(i) == tid and n.kind == "cat":
                return n.shape[0]
        return nodes[x].shape[0]

    # -- execution --------------------------------------------------------------
    def alloc_input(self, batch: int = None):
        """Zeroed flat bf16 input buffer in the layout ``wsi_preproc_run`` writes (borders stay zero).

        Always sized for ``max_batch`` (the parity-plane layout fixes the plane stride) plus read slack."""
        import torch

        return torch.zeros(self.input_bytes // 2, dtype=torch.bfloat16, device=self.device)

    def pack_input(self, x_nchw):
        """Test helper: fp32/bf16 ``(B, 3, S, S)`` tensor -> input buffer in this network's layout
        (what the preprocess kernel would have written for already-normalised pixels)."""
        import torch

        lay = self.input_layout
        b, _, s, _ = x_nchw.shape
        hp, pt, pl = lay["out_H"], lay["pad_top"], lay["pad_left"]
        img = torch.zeros((self.max_batch, hp, hp, lay["channels"]), dtype=torch.bfloat16, device=self.device)
        img[:b, pt:pt + s, pl:pl + s, :3] = x_nchw.permute(0, 2, 3, 1).to(torch.bfloat16).to(self.device)
        buf = self.alloc_input()
        if lay["kind"] == "parity":
            planes = torch.stack([img[:, 0::2], img[:, 1::2]], 0)      # [2][maxB][P][2P][4]
            buf[: planes.numel()] = planes.reshape(-1)
        else:
            buf[: img.numel()] = img.reshape(-1)
        return buf

    def forward_into(self, x, probs, stream=None, batch=None):
        """Enqueue the forward of the first ``batch`` (default ``len(probs)``) images of input buffer ``x``
        (layout of :meth:`alloc_input`), writing fp32 probs [B][K]."""
        import torch

        b = int(probs.shape[0]) if batch is None else int(batch)
        assert x.dtype == torch.bfloat16 and x.is_contiguous() and x.device == self.device
        assert x.numel() * 2 >= self.input_bytes, "allocate the input with alloc_input()"
        assert probs.dtype == torch.float32 and probs.is_contiguous() and probs.numel() >= b * self.num_classes
        with torch.cuda.device(self.device):
            _cabi.call("wsi_net_forward", self._h, x.data_ptr(), b, probs.data_ptr(), _cabi.stream_ptr(stream))

    def __call__(self, x, batch: int):
        import torch

        probs = torch.empty((int(batch), self.num_classes), dtype=torch.float32, device=self.device)
        self.forward_into(x, probs)
        return probs.squeeze(1) if self.num_classes == 1 else probs

    def activation(self, node_index: int, batch: int):
        """Debug: bf16 activation tensor [batch][H][W][C] of IR node ``node_index`` after a forward."""
        import torch

        ptr, hwc = C.c_void_p(), (C.c_int * 3)()
        _cabi.call("wsi_net_tensor_ptr", self._h, self._tensor_of[node_index], C.byref(ptr), hwc)
        h, w, c = int(hwc[0]), int(hwc[1]), int(hwc[2])
        out = torch.empty((batch, h, w, c), dtype=torch.bfloat16, device=self.device)
        with torch.cuda.device(self.device):
            _cabi.call("wsi_net_read_tensor", self._h, self._tensor_of[node_index], int(batch), out.data_ptr(),
                       _cabi.stream_ptr())
        return out

    def channel_offset(self, node_index: int) -> int:
        """Channel offset of IR node ``node_index`` inside its (possibly concat) tensor."""
        return int(self._c_off.get(node_index, 0))

    def set_profiling(self, slots: int):
        _cabi.call("wsi_net_set_profiling", self._h, int(slots))

    def profile(self):
        """Per-op event timings: list of dict(kind, ms_total, flops_per_image) + number of forwards."""
        nmax = 512
        ms, fl, kd, nf = (C.c_double * nmax)(), (C.c_double * nmax)(), (C.c_int * nmax)(), C.c_int()
        _cabi.call("wsi_net_profile_read", self._h, ms, fl, kd, nmax, C.byref(nf))
        kinds = {0: "conv", 1: "maxpool", 2: "avgpool", 3: "head"}
        ops = [dict(kind=kinds[int(kd[i])], ms_total=float(ms[i]), flops_per_image=float(fl[i]))
               for i in range(self.num_ops)]
        return ops, int(nf.value)

    def close(self):
        if self._h:
            _cabi.load().wsi_net_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
