"""Drop-in for the reference's quantization/quantizer.py, backed by libbdq.so (hand-written sm_100a kernels).

Same names, signatures, defaults, return shapes/dtypes and exception types as the reference:
Round (quantizer.py:12-22), pseudo_quantize_tensor (:25-69), real_quantize_model_weight (:73-108),
pseudo_quantize_n2f3_tensor (:113-116), SteInt3AsymQuantizer (:119-151), SteInt2AsymQuantizer (:153-182),
SteN2F3Quantizer (:184-242).  Differences: the per-call NaN asserts (3 host syncs per call in the reference)
are opt-in via BDQ_CHECK_NAN=1, and tensors must live on a CUDA device (no CPU fallback).
"""
import gc
import os

import torch
import torch.nn as nn
from torch.autograd import Function

from . import _lib as L


def _check_nan(*ts):
    if os.environ.get("BDQ_CHECK_NAN", "0") == "1":
        for t in ts:
            assert torch.isnan(t).sum() == 0


def _group(shape, q_group_size):
    if q_group_size > 0:
        assert shape[-1] % q_group_size == 0
        return q_group_size
    elif q_group_size == -1:
        return shape[-1]
    raise AssertionError("q_group_size must be > 0 or -1")


def int_fakequant(w, n_bit, group, round_mode, out=None, want_scale_zp=False, want_codes=False,
                  clip_min=None, clip_max=None):
    """Thin wrapper over bdq_int_fakequant_fwd.  Returns (dq, scales|None, zeros|None, codes|None)."""
    L.require_cuda(w, clip_min, clip_max)
    w = w.contiguous()
    n = w.numel()
    dq = torch.empty_like(w) if out is None else out
    ng = n // group
    scales = torch.empty(ng, dtype=w.dtype, device=w.device) if want_scale_zp else None
    zeros = torch.empty(ng, dtype=w.dtype, device=w.device) if want_scale_zp else None
    codes = torch.empty(w.shape, dtype=torch.uint8, device=w.device) if want_codes else None
    if clip_min is not None:
        clip_min = clip_min.to(w.dtype).contiguous()
        clip_max = clip_max.to(w.dtype).contiguous()
        assert clip_min.numel() == ng and clip_max.numel() == ng
    with L.on(w.device):
        rc = L.lib().bdq_int_fakequant_fwd(L.ptr(w), L.ptr(dq), L.ptr(scales), L.ptr(zeros), L.ptr(codes),
                                           L.ptr(clip_min), L.ptr(clip_max), L.i64(n), int(group), int(n_bit),
                                           L.dtype_code(w), int(round_mode), L.semantics(), L.stream_ptr(w.device))
    L.check(rc, "bdq_int_fakequant_fwd")
    return dq, scales, zeros, codes


def nf3_fakequant(w, group, want_codes=False, want_scales=False, clip_min=None, clip_max=None):
    """Thin wrapper over bdq_nf3_fakequant_fwd.  Returns (dq, scale_pos|None, scale_neg|None, codes|None)."""
    L.require_cuda(w, clip_min, clip_max)
    w = w.contiguous()
    n = w.numel()
    ng = n // group
    dq = torch.empty_like(w)
    sp = torch.empty(ng, dtype=w.dtype, device=w.device) if want_scales else None
    sn = torch.empty(ng, dtype=w.dtype, device=w.device) if want_scales else None
    codes = torch.empty(w.shape, dtype=torch.uint8, device=w.device) if want_codes else None
    if clip_min is not None:
        clip_min = clip_min.to(w.dtype).contiguous()
        clip_max = clip_max.to(w.dtype).contiguous()
        assert clip_min.numel() == ng and clip_max.numel() == ng
    with L.on(w.device):
        rc = L.lib().bdq_nf3_fakequant_fwd(L.ptr(w), L.ptr(dq), L.ptr(sp), L.ptr(sn), L.ptr(codes), L.ptr(clip_min),
                                           L.ptr(clip_max), L.i64(n), int(group), L.dtype_code(w), L.semantics(),
                                           L.stream_ptr(w.device))
    L.check(rc, "bdq_nf3_fakequant_fwd")
    return dq, sp, sn, codes


class Round(Function):
    """sign(x) * floor(|x| + 0.5) with identity backward (quantizer.py:12-22).  Elementwise; kept in torch ops
    because the fused quantizer kernels never materialise it -- it is API surface only."""

    @staticmethod
    def forward(self, input):
        sign = torch.sign(input)
        output = sign * torch.floor(torch.abs(input) + 0.5)
        return output

    @staticmethod
    def backward(self, grad_output):
        grad_input = grad_output.clone()
        return grad_input


def pseudo_quantize_tensor(w, n_bit=8, zero_point=True, q_group_size=-1, inplace=False, get_scale_zp=False):
    """quantizer.py:25-69.  One fused kernel instead of ~14 eager passes."""
    org_w_shape = w.shape
    group = _group(org_w_shape, q_group_size)
    if not zero_point:
        # the reference's zero_point=False branch dereferences an unbound local (quantizer.py:45) and never ran
        raise UnboundLocalError("cannot access local variable 'min_val' where it is not associated with a value")
    _check_nan(w)
    src = w if w.is_contiguous() else w.contiguous()
    # inplace: the kernels read w through ld.global.nc / __restrict__ (read-only contract), so they never write the buffer they
    # read -- the result goes to a temporary and is copied back below (the reference's in-place ops allocate temporaries too)
    dq, scales, zeros, _ = int_fakequant(src, n_bit, group, L.ROUND_HALF_EVEN, out=None, want_scale_zp=get_scale_zp)
    if inplace and dq is not w:
        w.copy_(dq.reshape(org_w_shape))
        dq = w
    _check_nan(dq)
    dq = dq.reshape(org_w_shape)
    if get_scale_zp:
        return dq, scales.view(dq.shape[0], -1), zeros.view(dq.shape[0], -1)
    return dq


@torch.no_grad()
def real_quantize_model_weight(model, w_bit, q_config, init_only=False):
    """quantizer.py:73-108: replace every decoder-block nn.Linear by a packed WQLinear."""
    from .pre_quant import get_blocks, get_named_linears, set_op_by_name
    from .qmodule import WQLinear

    assert q_config["zero_point"], "We only support zero_point quantization now."
    layers = get_blocks(model)
    for i in range(len(layers)):
        layer = layers[i]
        named_linears = get_named_linears(layer)
        for name, module in named_linears.items():
            if init_only:
                q_linear = WQLinear.from_linear(module, w_bit, q_config["q_group_size"], True)
                q_linear.to(next(layer.parameters()).device)
                set_op_by_name(layer, name, q_linear)
            else:
                module.cuda()
                qc = {k: v for k, v in q_config.items() if k in ("zero_point", "q_group_size")}
                module.weight.data, scales, zeros = pseudo_quantize_tensor(module.weight.data, n_bit=w_bit,
                                                                           get_scale_zp=True, **qc)
                q_linear = WQLinear.from_linear(module, w_bit, q_config["q_group_size"], False, scales, zeros)
                module.cpu()
                q_linear.to(next(layer.parameters()).device)
                set_op_by_name(layer, name, q_linear)
    torch.cuda.empty_cache()
    gc.collect()


class _SteIntFn(Function):
    """Fused forward (bdq_int_fakequant_fwd, half-away rounding) + fused backward (bdq_int_ste_bwd).
    Saves only the weight; the backward recomputes scale / zero / mask in registers."""

    @staticmethod
    def forward(ctx, x, group, bits):
        xc = x.contiguous()
        ctx.save_for_backward(xc)
        ctx.group, ctx.bits = group, bits
        dq, _, _, _ = int_fakequant(xc, bits, group, L.ROUND_HALF_AWAY)
        return dq

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        gy = gy.to(x.dtype).contiguous()
        L.require_cuda(gy)
        gw = torch.empty_like(x)
        with L.on(x.device):
            rc = L.lib().bdq_int_ste_bwd(L.ptr(x), L.ptr(gy), L.ptr(gw), L.i64(x.numel()), int(ctx.group), int(ctx.bits),
                                         L.dtype_code(x), L.semantics(), L.stream_ptr(x.device))
        L.check(rc, "bdq_int_ste_bwd")
        return gw, None, None


class _SteNf3Fn(Function):
    @staticmethod
    def forward(ctx, x, group):
        xc = x.contiguous()
        ctx.save_for_backward(xc)
        ctx.group = group
        dq, _, _, _ = nf3_fakequant(xc, group)
        return dq

    @staticmethod
    def backward(ctx, gy):
        (x,) = ctx.saved_tensors
        gy = gy.to(x.dtype).contiguous()
        L.require_cuda(gy)
        gw = torch.empty_like(x)
        with L.on(x.device):
            rc = L.lib().bdq_nf3_ste_bwd(L.ptr(x), L.ptr(gy), L.ptr(gw), L.i64(x.numel()), int(ctx.group),
                                         L.dtype_code(x), L.semantics(), L.stream_ptr(x.device))
        L.check(rc, "bdq_nf3_ste_bwd")
        return gw, None


def pseudo_quantize_n2f3_tensor(w, q_group_size=-1):
    """quantizer.py:113-116."""
    quantizer = SteN2F3Quantizer(q_group_size=q_group_size)
    w = quantizer(w)
    return w


class SteInt3AsymQuantizer(nn.Module):
    """quantizer.py:119-151."""

    def __init__(self, q_group_size=128):
        super().__init__()
        self.q_group_size = q_group_size
        self.bit = 3

    def forward(self, x):
        org_w_shape = x.shape
        group = _group(org_w_shape, self.q_group_size)
        _check_nan(x)
        return _SteIntFn.apply(x, group, self.bit).reshape(org_w_shape)


class SteInt2AsymQuantizer(nn.Module):
    """quantizer.py:153-182."""

    def __init__(self, q_group_size=64):
        super().__init__()
        self.q_group_size = q_group_size
        self.bit = 2

    def forward(self, x):
        org_w_shape = x.shape
        if self.q_group_size > 0:
            assert org_w_shape[-1] % self.q_group_size == 0
            group = self.q_group_size
        else:
            assert x.dim() == 2  # the reference only reshapes for q_group_size > 0 (quantizer.py:161-164)
            group = org_w_shape[-1]
        _check_nan(x)
        return _SteIntFn.apply(x, group, self.bit).reshape(org_w_shape)


class SteN2F3Quantizer(nn.Module):
    """quantizer.py:184-242."""

    def __init__(self, q_group_size=128):
        super().__init__()
        self.q_group_size = q_group_size

    def forward(self, x):
        org_w_shape = x.shape
        group = _group(org_w_shape, self.q_group_size)
        return _SteNf3Fn.apply(x, group).reshape(org_w_shape)
