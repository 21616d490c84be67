"""Helpers for the GPU network tests: drive single layers through the raw C ABI."""

import ctypes as C

import numpy as np
import torch

from wsinsight_b200 import _cabi


def bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).float()


def run_single_conv(x_nchw, w, b, stride, pad, relu=False, residual=None, max_batch=None, expect_flat=None):
    """x: fp32 NCHW (values already bf16-representable).  Returns fp32 NCHW output of the B200 conv."""
    lib = _cabi.load()
    B, Cin, H, W = x_nchw.shape
    Cout, _, KH, KW = w.shape
    ph, pw = (pad, pad) if isinstance(pad, int) else pad
    Ho, Wo = (H + 2 * ph - KH) // stride + 1, (W + 2 * pw - KW) // stride + 1
    net = C.c_void_p()
    _cabi.call("wsi_net_create", int(max_batch or B), C.byref(net))
    try:
        tin, tout, tres = C.c_int(), C.c_int(), C.c_int(-1)
        # the input is an internal tensor filled through wsi_net_write_tensor, so that eligible shapes
        # take the flat-slab kernel exactly as they do inside a network
        _cabi.call("wsi_net_add_tensor", net, H, W, Cin, 0, C.byref(tin))
        _cabi.call("wsi_net_add_tensor", net, Ho, Wo, Cout, 0, C.byref(tout))
        res_id = -1
        if residual is not None:
            _cabi.call("wsi_net_add_tensor", net, Ho, Wo, Cout, 0, C.byref(tres))
            res_id = tres.value
        wc = w.contiguous().float().cpu()
        bc = b.contiguous().float().cpu() if b is not None else None
        _cabi.call("wsi_net_add_conv", net, tin.value, tout.value, 0, C.cast(wc.data_ptr(), _cabi.c_float_p),
                   C.cast(bc.data_ptr(), _cabi.c_float_p) if bc is not None else None, Cin, Cout, KH, KW, stride, ph, pw,
                   1 if relu else 0, res_id)
        _cabi.call("wsi_net_finalize", net)
        xd = x_nchw.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
        _cabi.call("wsi_net_write_tensor", net, tin.value, B, xd.data_ptr(), _cabi.stream_ptr())
        if residual is not None:
            rd = residual.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
            _cabi.call("wsi_net_write_tensor", net, res_id, B, rd.data_ptr(), _cabi.stream_ptr())
        _cabi.call("wsi_net_forward", net, None, B, None, _cabi.stream_ptr())
        out = torch.empty((B, Ho, Wo, Cout), dtype=torch.bfloat16, device="cuda")
        _cabi.call("wsi_net_read_tensor", net, tout.value, B, out.data_ptr(), _cabi.stream_ptr())
        if expect_flat is not None:
            assert bool(lib.wsi_net_op_is_flat(net, 0)) == expect_flat
        torch.cuda.synchronize()
        return out.float().permute(0, 3, 1, 2).cpu()
    finally:
        lib.wsi_net_destroy(net)


def conv_reference(x, w, b, stride, pad, relu=False, residual=None):
    y = torch.nn.functional.conv2d(x.double(), w.double(), None if b is None else b.double(), stride=stride, padding=pad)
    if residual is not None:
        y = y + residual.double()
    if relu:
        y = torch.relu(y)
    return y.float()


def emulate_ir(nodes, x_nchw, accum=torch.float64):
    """CPU emulation of the B200 network numerics: bf16 weights/activations, fp32 accumulate.

    Runs the lowered IR (wsinsight_b200.modellib.lowering) node by node in torch and rounds every
    stored activation to bf16, exactly where the kernels do.  Returns ({node index: fp32 NCHW
    activation}, probabilities).  Differences to the GPU are accumulation-order noise only; that
    noise is amplified layer by layer (a bf16 rounding flip is 2^-8 relative), so callers calibrate
    their tolerance by running the emulation twice, with ``accum`` fp64 and fp32.
    """
    import torch.nn.functional as F

    acts = {0: bf16_round(x_nchw)}
    probs = None
    from wsinsight_b200.modellib.engine import _topo_order

    for i in _topo_order(nodes):
        n = nodes[i]
        if n.kind == "conv":
            y = F.conv2d(acts[n.inputs[0]].to(accum), bf16_round(n.weight).to(accum),
                         None if n.bias is None else n.bias.to(accum), stride=n.stride, padding=tuple(n.pad))
            if n.residual is not None:
                y = y + acts[n.residual].to(accum)
            if n.relu:
                y = torch.relu(y)
            acts[i] = bf16_round(y.float())
        elif n.kind == "pool":
            a = acts[n.inputs[0]]
            if n.pool_max:
                y = F.max_pool2d(a, n.k, n.pstride, n.ppad)
            else:
                y = F.avg_pool2d(a, n.k, n.pstride, n.ppad, count_include_pad=n.count_include_pad)
            acts[i] = bf16_round(y)
        elif n.kind == "cat":
            acts[i] = torch.cat([acts[p] for p in n.inputs], 1)
        elif n.kind == "head":
            feat = acts[n.inputs[0]].double().mean((2, 3))
            logits = feat @ n.fc_w.double().t()
            if n.fc_b is not None:
                logits = logits + n.fc_b.double()
            logits = logits.float()
            probs = torch.softmax(logits, 1) if logits.shape[1] > 1 else torch.sigmoid(logits.squeeze(1))
    return acts, probs


def run_single_pool(x_nchw, is_max, k, stride, pad, count_include_pad=False, out_channels=None, c_off=0):
    """One pooling op through the raw C ABI; the output may be a channel slice of a wider (concat) tensor."""
    lib = _cabi.load()
    B, Cc, H, W = x_nchw.shape
    Ho, Wo = (H + 2 * pad - k) // stride + 1, (W + 2 * pad - k) // stride + 1
    Ct = int(out_channels or Cc)
    net = C.c_void_p()
    _cabi.call("wsi_net_create", B, C.byref(net))
    try:
        tin, tout = C.c_int(), C.c_int()
        _cabi.call("wsi_net_add_tensor", net, H, W, Cc, 0, C.byref(tin))
        _cabi.call("wsi_net_add_tensor", net, Ho, Wo, Ct, 0, C.byref(tout))
        _cabi.call("wsi_net_add_pool", net, tin.value, tout.value, int(c_off), 1 if is_max else 0, k, stride, pad,
                   1 if count_include_pad else 0)
        _cabi.call("wsi_net_finalize", net)
        xd = x_nchw.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
        _cabi.call("wsi_net_write_tensor", net, tin.value, B, xd.data_ptr(), _cabi.stream_ptr())
        _cabi.call("wsi_net_forward", net, None, B, None, _cabi.stream_ptr())
        out = torch.empty((B, Ho, Wo, Ct), dtype=torch.bfloat16, device="cuda")
        _cabi.call("wsi_net_read_tensor", net, tout.value, B, out.data_ptr(), _cabi.stream_ptr())
        torch.cuda.synchronize()
        return out.float().permute(0, 3, 1, 2).cpu()
    finally:
        lib.wsi_net_destroy(net)


def run_pool_bias(x_nchw, in_c_off, Cs, bias, relu, count_include_pad=False, out_channels=None, c_off=0):
    """3x3/1/1 average pool + bias + ReLU over channels [in_c_off, in_c_off + Cs) of the input (wsi_net_add_pool_bias)."""
    lib = _cabi.load()
    B, Cc, H, W = x_nchw.shape
    Ct = int(out_channels or Cs)
    net = C.c_void_p()
    _cabi.call("wsi_net_create", B, C.byref(net))
    try:
        tin, tout = C.c_int(), C.c_int()
        _cabi.call("wsi_net_add_tensor", net, H, W, Cc, 0, C.byref(tin))
        _cabi.call("wsi_net_add_tensor", net, H, W, Ct, 0, C.byref(tout))
        b = bias.contiguous().float()
        _cabi.call("wsi_net_add_pool_bias", net, tin.value, int(in_c_off), int(Cs), tout.value, int(c_off), 3, 1, 1,
                   1 if count_include_pad else 0, C.cast(b.data_ptr(), _cabi.c_float_p), 1 if relu else 0)
        _cabi.call("wsi_net_finalize", net)
        xd = x_nchw.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
        _cabi.call("wsi_net_write_tensor", net, tin.value, B, xd.data_ptr(), _cabi.stream_ptr())
        _cabi.call("wsi_net_forward", net, None, B, None, _cabi.stream_ptr())
        out = torch.empty((B, H, W, Ct), dtype=torch.bfloat16, device="cuda")
        _cabi.call("wsi_net_read_tensor", net, tout.value, B, out.data_ptr(), _cabi.stream_ptr())
        torch.cuda.synchronize()
        return out.float().permute(0, 3, 1, 2).cpu()
    finally:
        lib.wsi_net_destroy(net)


def run_conv_slice(x_nchw, in_c_off, w, b, stride, pad, relu_cols=None, expect_flat=None):
    """Conv over channels [in_c_off, in_c_off + Cin) of a wider tensor (wsi_net_add_conv_slice) with ReLU on the first
    ``relu_cols`` output channels only (wsi_net_set_conv_relu_cols; None = plain ReLU on all)."""
    lib = _cabi.load()
    B, Ctot, H, W = x_nchw.shape
    Cout, Cin, KH, KW = w.shape
    Ho, Wo = (H + 2 * pad - KH) // stride + 1, (W + 2 * pad - KW) // stride + 1
    net = C.c_void_p()
    _cabi.call("wsi_net_create", B, C.byref(net))
    try:
        tin, tout = C.c_int(), C.c_int()
        _cabi.call("wsi_net_add_tensor", net, H, W, Ctot, 0, C.byref(tin))
        _cabi.call("wsi_net_add_tensor", net, Ho, Wo, Cout, 0, C.byref(tout))
        wc, bc = w.contiguous().float().cpu(), b.contiguous().float().cpu()
        _cabi.call("wsi_net_add_conv_slice", net, tin.value, int(in_c_off), tout.value, 0,
                   C.cast(wc.data_ptr(), _cabi.c_float_p), C.cast(bc.data_ptr(), _cabi.c_float_p), Cin, Cout, KH, KW,
                   stride, pad, pad, 1, -1)
        if relu_cols is not None:
            _cabi.call("wsi_net_set_conv_relu_cols", net, int(relu_cols))
        _cabi.call("wsi_net_finalize", net)
        xd = x_nchw.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
        _cabi.call("wsi_net_write_tensor", net, tin.value, B, xd.data_ptr(), _cabi.stream_ptr())
        _cabi.call("wsi_net_forward", net, None, B, None, _cabi.stream_ptr())
        out = torch.empty((B, Ho, Wo, Cout), dtype=torch.bfloat16, device="cuda")
        _cabi.call("wsi_net_read_tensor", net, tout.value, B, out.data_ptr(), _cabi.stream_ptr())
        if expect_flat is not None:
            assert bool(lib.wsi_net_op_is_flat(net, 0)) == expect_flat
        torch.cuda.synchronize()
        return out.float().permute(0, 3, 1, 2).cpu()
    finally:
        lib.wsi_net_destroy(net)
