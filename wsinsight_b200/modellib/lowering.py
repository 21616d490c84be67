"""TorchScript graph -> layer list for the B200 network runtime (csrc/net.cu).

Zoo model files are ``torch.jit.script``/``trace`` + ``torch.jit.freeze`` outputs
(reference ``scripts/convert_to_torchscript.py:38-50``): weights are graph constants with
BatchNorm already folded.  This module walks the inlined graph of such a file
(freezing it first if it is not frozen) and produces a small IR:

    Conv(in, w, b, stride, pad, relu, residual)   Pool(in, kind, k, s, p, count_include_pad)
    Cat(inputs)                                   Head(in, fc_w, fc_b)   [GAP + linear + softmax]

ReLU and residual adds are fused into the producing convolution.  Anything else raises
:class:`UnknownArchitectureError` -- there is deliberately no fallback to running the
TorchScript module itself.
"""

from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional

from ..errors import UnknownArchitectureError


@dataclasses.dataclass
class IRNode:
    kind: str                      # "input" | "conv" | "pool" | "cat" | "head"
    inputs: List[int]
    shape: tuple = ()              # (C, H, W) of the output
    # conv
    weight: object = None          # torch fp32 OIHW
    bias: object = None            # torch fp32 (O,) or None
    stride: int = 1
    pad: tuple = (0, 0)
    relu: bool = False
    residual: Optional[int] = None
    # pool
    pool_max: bool = True
    k: int = 0
    pstride: int = 0
    ppad: int = 0
    count_include_pad: bool = True
    # head
    fc_w: object = None
    fc_b: object = None
    n_uses: int = 0


def _pair(v):
    v = list(v) if isinstance(v, (list, tuple)) else [v, v]
    if len(v) == 1:
        v = [v[0], v[0]]
    return int(v[0]), int(v[1])


def _const(value):
    node = value.node()
    if node.kind() == "prim::ListConstruct":      # serialised graphs rebuild constant lists
        return [_const(i) for i in node.inputs()]
    if node.kind() != "prim::Constant":
        raise UnknownArchitectureError(f"expected a constant for {value.debugName()}, got {node.kind()}")
    return value.toIValue()


def load_frozen(path_or_module):
    """torch.jit.load + (if needed) eval + freeze, always on the CPU."""
    import torch

    m = path_or_module
    if not isinstance(m, torch.jit.ScriptModule):
        m = torch.jit.load(str(path_or_module), map_location="cpu")
    m = m.eval()
    has_params = any(True for _ in m.parameters())
    if has_params:
        m = torch.jit.freeze(m)
    return m


def lower_torchscript(path_or_module, input_hw: int) -> List[IRNode]:
    """Lower a classification CNN to the IR.  ``input_hw``: S of the (3, S, S) model input."""
    import torch

    m = load_frozen(path_or_module)
    g = m.graph
    torch._C._jit_pass_inline(g)
    nodes: List[IRNode] = [IRNode("input", [], (3, input_hw, input_hw))]
    env: Dict[str, int] = {}
    lists: Dict[str, List[int]] = {}
    gi = list(g.inputs())
    tensor_inputs = [v for v in gi if str(v.type()) == "Tensor"]
    if len(tensor_inputs) != 1:
        raise UnknownArchitectureError(f"expected exactly one tensor input, got {len(tensor_inputs)}")
    env[tensor_inputs[0].debugName()] = 0
    pending_gap: Dict[str, int] = {}   # value produced by global-average-pool(+flatten) -> source node
    out_node = None

    def src(v):
        name = v.debugName()
        if name not in env:
            raise UnknownArchitectureError(f"value {name} is produced by an unsupported op ({v.node().kind()})")
        return env[name]

    for n in g.nodes():
        kind = n.kind()
        ins = list(n.inputs())
        outs = list(n.outputs())
        if kind == "prim::Constant":
            continue
        if kind == "prim::ListConstruct":
            if all(str(i.type()) == "Tensor" for i in ins) and ins:
                lists[outs[0].debugName()] = [src(i) for i in ins]
            continue
        if not outs or all(str(o.type()) != "Tensor" for o in outs):
            continue   # shape bookkeeping (aten::size/len/gt, prim::If without outputs, ...)
        oname = outs[0].debugName()
        if kind in ("aten::conv2d", "aten::_convolution"):
            x = src(ins[0])
            w = _const(ins[1]).detach().float().contiguous()
            b = _const(ins[2])
            b = None if b is None else b.detach().float().contiguous()
            stride, pad, dil = _pair(_const(ins[3])), _pair(_const(ins[4])), _pair(_const(ins[5]))
            groups = int(_const(ins[8] if kind == "aten::_convolution" else ins[6]))
            if kind == "aten::_convolution" and bool(_const(ins[6])):
                raise UnknownArchitectureError("transposed convolution is not supported")
            if dil != (1, 1) or groups != 1 or stride[0] != stride[1]:
                raise UnknownArchitectureError(f"conv with dilation {dil} / groups {groups} / stride {stride}")
            cin, h, wd = nodes[x].shape
            if w.shape[1] != cin:
                raise UnknownArchitectureError(f"conv expects {w.shape[1]} input channels, tensor has {cin}")
            ho = (h + 2 * pad[0] - w.shape[2]) // stride[0] + 1
            wo = (wd + 2 * pad[1] - w.shape[3]) // stride[0] + 1
            nodes.append(IRNode("conv", [x], (int(w.shape[0]), ho, wo), weight=w, bias=b, stride=stride[0],
                                pad=pad))
            env[oname] = len(nodes) - 1
        elif kind in ("aten::relu", "aten::relu_"):
            x = src(ins[0])
            if nodes[x].kind != "conv" or nodes[x].relu:
                raise UnknownArchitectureError("ReLU that does not follow a convolution / residual add")
            if len(ins[0].uses()) != 1:
                # the ReLU is fused into the producing conv: a second reader of the pre-activation value
                # (pre-activation blocks, a raw branch) would silently see the activated tensor
                raise UnknownArchitectureError("ReLU input is also read un-activated by another op")
            nodes[x].relu = True
            env[oname] = x
        elif kind in ("aten::add", "aten::add_"):
            if float(_const(ins[2])) != 1.0:
                raise UnknownArchitectureError("add with alpha != 1")
            a, b = src(ins[0]), src(ins[1])

            def fusable(j, v):
                # the add is fused into the conv that produced `v`: nothing else may read the pre-add value
                return nodes[j].kind == "conv" and not nodes[j].relu and nodes[j].residual is None \
                    and len(v.uses()) == 1

            pick = next(((t_, o_) for t_, v_, o_ in ((a, ins[0], b), (b, ins[1], a)) if t_ != o_ and fusable(t_, v_)), None)
            if pick is None or nodes[pick[0]].shape != nodes[pick[1]].shape:
                raise UnknownArchitectureError("residual add that cannot be fused into a convolution")
            tgt, other = pick
            t = nodes[tgt]
            t.residual = other
            env[oname] = tgt
        elif kind == "aten::max_pool2d":
            x = src(ins[0])
            k, s, p, d = _pair(_const(ins[1])), _const(ins[2]), _pair(_const(ins[3])), _pair(_const(ins[4]))
            s = _pair(s) if s else k
            if bool(_const(ins[5])) or d != (1, 1) or k[0] != k[1] or s[0] != s[1] or p[0] != p[1]:
                raise UnknownArchitectureError("max_pool2d variant (ceil_mode/dilation/non-square)")
            c, h, wd = nodes[x].shape
            nodes.append(IRNode("pool", [x], (c, (h + 2 * p[0] - k[0]) // s[0] + 1, (wd + 2 * p[0] - k[0]) // s[0] + 1),
                                pool_max=True, k=k[0], pstride=s[0], ppad=p[0]))
            env[oname] = len(nodes) - 1
        elif kind == "aten::avg_pool2d":
            x = src(ins[0])
            k, s, p = _pair(_const(ins[1])), _const(ins[2]), _pair(_const(ins[3]))
            s = _pair(s) if s else k
            ceil_mode, cip, div = bool(_const(ins[4])), bool(_const(ins[5])), _const(ins[6])
            c, h, wd = nodes[x].shape
            if ceil_mode or div is not None or k[0] != k[1] or s[0] != s[1] or p[0] != p[1]:
                raise UnknownArchitectureError("avg_pool2d variant")
            if k[0] == h and k[1] == wd and p == (0, 0):
                pending_gap[oname] = x       # global average pool
                continue
            nodes.append(IRNode("pool", [x], (c, (h + 2 * p[0] - k[0]) // s[0] + 1, (wd + 2 * p[0] - k[0]) // s[0] + 1),
                                pool_max=False, k=k[0], pstride=s[0], ppad=p[0], count_include_pad=cip))
            env[oname] = len(nodes) - 1
        elif kind == "aten::adaptive_avg_pool2d":
            if _pair(_const(ins[1])) != (1, 1):
                raise UnknownArchitectureError("adaptive_avg_pool2d to a size other than 1x1")
            pending_gap[oname] = src(ins[0])
        elif kind in ("aten::flatten", "aten::view", "aten::reshape", "aten::dropout", "aten::dropout_",
                      "aten::feature_dropout", "aten::contiguous"):
            iname = ins[0].debugName()
            if iname in pending_gap:
                pending_gap[oname] = pending_gap[iname]
            elif kind.startswith("aten::dropout") or kind in ("aten::feature_dropout", "aten::contiguous"):
                env[oname] = src(ins[0])
            else:
                raise UnknownArchitectureError(f"{kind} outside the classifier tail")
        elif kind == "aten::cat":
            if int(_const(ins[1])) != 1:
                raise UnknownArchitectureError("cat along a dimension other than channels")
            parts = lists.get(ins[0].debugName())
            if not parts:
                raise UnknownArchitectureError("cat of a non-literal list")
            h, wd = nodes[parts[0]].shape[1:]
            if any(nodes[p].shape[1:] != (h, wd) for p in parts):
                raise UnknownArchitectureError("cat of tensors with different spatial sizes")
            nodes.append(IRNode("cat", list(parts), (sum(nodes[p].shape[0] for p in parts), h, wd)))
            env[oname] = len(nodes) - 1
        elif kind in ("aten::linear", "aten::addmm"):
            if kind == "aten::linear":
                xin, w, b = ins[0], _const(ins[1]), _const(ins[2])
            else:   # addmm(bias, x, w.t())
                raise UnknownArchitectureError("aten::addmm classifier (trace with aten::linear instead)")
            iname = xin.debugName()
            if iname not in pending_gap:
                raise UnknownArchitectureError("linear layer that does not follow global average pooling")
            x = pending_gap[iname]
            w = w.detach().float().contiguous()
            if w.shape[1] != nodes[x].shape[0]:
                raise UnknownArchitectureError("classifier width does not match the pooled features")
            nodes.append(IRNode("head", [x], (int(w.shape[0]), 1, 1), fc_w=w,
                                fc_b=None if b is None else b.detach().float().contiguous()))
            env[oname] = len(nodes) - 1
            out_node = len(nodes) - 1
        else:
            raise UnknownArchitectureError(f"unsupported TorchScript op {kind}")

    ret = list(g.outputs())
    if out_node is None or len(ret) != 1 or env.get(ret[0].debugName()) != out_node:
        raise UnknownArchitectureError("graph does not end in global-average-pool -> linear")
    for nd in nodes:
        for i in nd.inputs:
            nodes[i].n_uses += 1
        if nd.residual is not None:
            nodes[nd.residual].n_uses += 1
    return nodes
