"""Packed NF3 weights (3 bits + two fp16 scales per group).  The reference stops at the NF3 fake-quantiser
(quantization/quantizer.py:184-242): it has no packed NF3 format, no NF3 WQLinear and no NF3 kernels, so this module is
this package's own extension of `qmodule.WQLinear` to NF3, with the layout documented in include/bdq.h.  Contract:
`dequant_nf3(*pack_nf3(w))` equals `pseudo_quantize_n2f3_tensor(w)` bit for bit in fp16.

First version: decode (<= 8 rows) runs `bdq_gemv_nf3` (CUDA-core table decode); longer inputs dequantise the packed
weight once per call (`bdq_nf3_dequant`) and use a dense fp16 matmul -- a library GEMM, not yet the tcgen05 path of the
int2 format."""
import torch
import torch.nn as nn

from . import _lib as L
from .quantizer import nf3_fakequant


@torch.no_grad()
def pack_nf3(w, group_size=128, clip_min=None, clip_max=None):
    """[OC, IC] fp16 weight -> (qweight int32 [OC, IC/16], qhigh int32 [OC, IC/32], scales fp16 [OC, 2*IC/G])."""
    L.require_cuda(w)
    if w.dtype != torch.float16:
        raise TypeError("pack_nf3 packs fp16 weights (the inference dtype of WQLinear)")
    oc, ic = w.shape
    if ic % group_size or group_size % 32:
        raise ValueError("in_features must be a multiple of group_size, group_size a multiple of 32")
    _, sp, sn, codes = nf3_fakequant(w, group_size, want_codes=True, want_scales=True, clip_min=clip_min, clip_max=clip_max)
    nan = (sp == 0) | (sn == 0)           # the reference's 0/0 groups are NaN throughout
    sp = torch.where(nan, torch.full_like(sp, float("nan")), sp)
    sn = torch.where(nan, torch.full_like(sn, float("nan")), sn)
    scales = torch.stack([sp.view(oc, -1), sn.view(oc, -1)], dim=-1).reshape(oc, -1).contiguous()
    qweight = torch.empty((oc, ic // 16), dtype=torch.int32, device=w.device)
    qhigh = torch.empty((oc, ic // 32), dtype=torch.int32, device=w.device)
    with L.on(w.device):
        rc = L.lib().bdq_pack_nf3(L.ptr(codes), L.ptr(qweight), L.ptr(qhigh), L.i64(oc), L.i64(ic), L.stream_ptr(w.device))
    L.check(rc, "bdq_pack_nf3")
    return qweight, qhigh, scales


@torch.no_grad()
def dequant_nf3(qweight, qhigh, scales, group_size=128):
    """Packed NF3 -> [OC, IC] fp16 (the fake-quantised weight)."""
    L.require_cuda(qweight, qhigh, scales)
    oc, ic = qweight.shape[0], qweight.shape[1] * 16
    out = torch.empty((oc, ic), dtype=torch.float16, device=qweight.device)
    with L.on(qweight.device):
        rc = L.lib().bdq_nf3_dequant(L.ptr(qweight), L.ptr(qhigh), L.ptr(scales), L.ptr(out), L.i64(oc), L.i64(ic),
                                     int(group_size), L.stream_ptr(qweight.device))
    L.check(rc, "bdq_nf3_dequant")
    return out


@torch.no_grad()
def gemv_nf3(x, qweight, qhigh, scales, group_size=128):
    """[B, IC] fp16 x packed NF3 [OC, IC] -> [B, OC] fp16."""
    L.require_cuda(x, qweight, qhigh, scales)
    if x.dtype != torch.float16:
        raise TypeError("gemv_nf3 takes fp16 activations")
    x = x.contiguous()
    B, ic = x.shape
    oc = qweight.shape[0]
    assert qweight.shape[1] * 16 == ic
    out = torch.empty((B, oc), dtype=torch.float16, device=x.device)
    with L.on(x.device):
        rc = L.lib().bdq_gemv_nf3(L.ptr(x), L.ptr(qweight), L.ptr(qhigh), L.ptr(scales), L.ptr(out), L.i64(B), L.i64(ic),
                                  L.i64(oc), int(group_size), L.stream_ptr(x.device))
    L.check(rc, "bdq_gemv_nf3")
    return out


def gemm_f16(x, w):
    """out [M, OC] = x [M, IC] @ w [OC, IC]^T on this library's tcgen05 GEMM (bdq_gemm_f16: both operands by TMA, fp32
    accumulate, fp16 out)."""
    L.require_cuda(x, w)
    if x.dtype != torch.float16 or w.dtype != torch.float16:
        raise TypeError("gemm_f16 expects float16 tensors")
    x, w = x.contiguous(), w.contiguous()
    M, ic = x.shape
    oc = w.shape[0]
    assert w.shape[1] == ic
    out = torch.empty((M, oc), dtype=torch.float16, device=x.device)
    with L.on(x.device):
        rc = L.lib().bdq_gemm_f16(L.ptr(x), L.ptr(w), L.ptr(out), L.i64(M), L.i64(ic), L.i64(oc), L.stream_ptr(x.device))
    L.check(rc, "bdq_gemm_f16")
    return out


class WQLinearNF3(nn.Module):
    """NF3 counterpart of qmodule.WQLinear: buffers `qweight`, `qhigh`, `scales`, `bias` are the on-disk format."""

    def __init__(self, group_size, in_features, out_features, bias, dev):
        super().__init__()
        if in_features % group_size or group_size % 32:
            raise ValueError("in_features must be a multiple of group_size, group_size a multiple of 32")
        self.in_features, self.out_features, self.group_size, self.w_bit = in_features, out_features, group_size, 3
        self.register_buffer("qweight", torch.zeros((out_features, in_features // 16), dtype=torch.int32, device=dev))
        self.register_buffer("qhigh", torch.zeros((out_features, in_features // 32), dtype=torch.int32, device=dev))
        self.register_buffer("scales", torch.zeros((out_features, 2 * in_features // group_size), dtype=torch.float16, device=dev))
        if bias:
            self.register_buffer("bias", torch.zeros(out_features, dtype=torch.float16, device=dev))
        else:
            self.bias = None

    @classmethod
    def from_linear(cls, linear, group_size=128, init_only=False):
        m = cls(group_size, linear.in_features, linear.out_features, linear.bias is not None, linear.weight.device)
        if init_only:
            return m
        m.qweight, m.qhigh, m.scales = pack_nf3(linear.weight.data.half(), group_size)
        if linear.bias is not None:
            m.bias = linear.bias.data.clone().half()
        return m

    @torch.no_grad()
    def forward(self, x):
        out_shape = x.shape[:-1] + (self.out_features,)
        x2 = x.reshape(-1, x.shape[-1])
        if x2.shape[0] <= 8:
            y = gemv_nf3(x2, self.qweight, self.qhigh, self.scales, self.group_size)
        else:
            # prefill: dequantise once (bdq_nf3_dequant, bit-equal to the NF3 fake-quantiser) and run this library's own tcgen05
            # GEMM on it -- no cuBLAS on the path.  The 3-bit table decode costs ~3x the int2 digit trick per weight, too much to
            # hide behind the MMAs inside the GEMM's producers; the dequantised copy costs 4 B/weight of HBM traffic per call.
            w16 = dequant_nf3(self.qweight, self.qhigh, self.scales, self.group_size)
            if self.in_features % 64 == 0 and self.out_features % 32 == 0:
                y = gemm_f16(x2.half(), w16)
            else:
                raise L.BdqError("WQLinearNF3 prefill needs in_features % 64 == 0 and out_features % 32 == 0")
        if self.bias is not None:
            y = y + self.bias
        return y.reshape(out_shape)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, w_bit=3 (nf3), group_size={self.group_size}"
