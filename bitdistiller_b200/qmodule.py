"""Drop-in for the reference's quantization/qmodule.py (WQLinear, calculate_zeros_width), backed by libbdq.so.

State-dict keys (`qweight`, `qzeros`, `scales`, `bias`) and buffer shapes are the reference's on-disk format
(qmodule.py:62-66).  Differences, all deliberate:
  * from_linear packs with one fused kernel (bdq_pack_wq) instead of in_features + 2*in_features/pack eager launches;
  * forward never re-transposes the packed tensors per call (the reference makes three .T.contiguous() copies of the
    whole layer for every prefill, qmodule.py:154-158); rows <= 8 go to the GEMV, more rows to the tcgen05 GEMM;
  * M is the true number of rows (the reference passes x.shape[-2], wrong for batch > 1).
"""
import torch
import torch.nn as nn

from . import _lib as L
from . import awq_inference_engine

USE_TRITON = False  # no Triton anywhere in this package; kept because callers import the name


def make_divisible(c, divisor):
    return (c + divisor - 1) // divisor


def calculate_zeros_width(in_features, group_size=128, pack_num=8):
    """qmodule.py:16-28."""
    if group_size >= 128:
        size_multiplier = 1
    elif group_size == 64:
        size_multiplier = 2
    elif group_size == 32:
        size_multiplier = 4
    else:
        raise NotImplementedError
    base_width = make_divisible(in_features // group_size, pack_num)
    base_width = make_divisible(base_width, size_multiplier) * size_multiplier
    return base_width


class ScaledActivation(nn.Module):
    """qmodule.py:31-38."""

    def __init__(self, module, scales):
        super().__init__()
        self.act = module
        self.scales = nn.Parameter(scales.data)

    def forward(self, x):
        return self.act(x) / self.scales.view(1, 1, -1).to(x.device)


def pack_wq(w_dq, scales, zeros, w_bit, group_size):
    """Thin wrapper over bdq_pack_wq -> (qweight int32 [out, in/pack], qzeros int32 [out, zw], qscales fp16 [out, zw*pack])."""
    L.require_cuda(w_dq, scales, zeros)
    rows, cols = w_dq.shape
    pack_num = 32 // w_bit
    G = group_size if group_size != -1 else cols
    zw = calculate_zeros_width(cols, G, pack_num)
    w_dq = w_dq.contiguous()
    scales = scales.to(w_dq.dtype).contiguous()
    zeros = zeros.to(w_dq.dtype).contiguous()
    assert scales.numel() == rows * (cols // G) and zeros.numel() == scales.numel()
    dev = w_dq.device
    qweight = torch.empty((rows, cols // pack_num), dtype=torch.int32, device=dev)
    qzeros = torch.empty((rows, zw), dtype=torch.int32, device=dev)
    qscales = torch.empty((rows, zw * pack_num), dtype=torch.float16, device=dev)
    with L.on(dev):
        rc = L.lib().bdq_pack_wq(L.ptr(w_dq), L.ptr(scales), L.ptr(zeros), L.ptr(qweight), L.ptr(qzeros), L.ptr(qscales),
                                 L.i64(rows), L.i64(cols), int(G), int(w_bit), L.dtype_code(w_dq), L.stream_ptr(dev))
    L.check(rc, "bdq_pack_wq")
    return qweight, qzeros, qscales


class WQLinear(nn.Module):
    """qmodule.py:41-178."""

    def __init__(self, w_bit, group_size, in_features, out_features, bias, dev):
        super().__init__()
        if w_bit not in [4, 2]:
            raise NotImplementedError("Only 4,2-bit are supported for now.")
        self.in_features = in_features
        self.out_features = out_features
        self.w_bit = w_bit
        self.offset = 0x0F if w_bit == 4 else 0x03
        self.group_size = group_size if group_size != -1 else in_features
        self.split_k_iters = pack_num = (32 // self.w_bit)
        assert self.in_features % self.group_size == 0
        assert out_features % (32 // self.w_bit) == 0
        zw = calculate_zeros_width(in_features, self.group_size, pack_num)
        self.register_buffer('qweight', torch.zeros((out_features, in_features // pack_num), dtype=torch.int32, device=dev))
        self.register_buffer('qzeros', torch.zeros((out_features, zw), dtype=torch.int32, device=dev))
        self.register_buffer('scales', torch.zeros((out_features, zw * pack_num), dtype=torch.float16, device=dev))
        if bias:
            self.register_buffer('bias', torch.zeros((out_features), dtype=torch.float16, device=dev))
        else:
            self.bias = None

    @classmethod
    def from_linear(cls, linear, w_bit, group_size, init_only=False, scales=None, zeros=None):
        awq_linear = cls(w_bit, group_size, linear.in_features, linear.out_features, linear.bias is not None,
                         linear.weight.device)
        if init_only:  # just prepare for loading sd
            return awq_linear
        # need scales and zeros info for real quantization
        assert scales is not None and zeros is not None
        qweight, qzeros, qscales = pack_wq(linear.weight.data, scales, zeros, awq_linear.w_bit, awq_linear.group_size)
        awq_linear.scales = qscales
        if linear.bias is not None:
            awq_linear.bias = linear.bias.clone().half()
        awq_linear.qweight = qweight
        awq_linear.qzeros = qzeros
        return awq_linear

    @torch.no_grad()
    def forward(self, x):
        out_shape = x.shape[:-1] + (self.out_features, )
        inputs = x.reshape(-1, x.shape[-1])
        if inputs.shape[0] > 8:
            out = awq_inference_engine.gemm_forward_cuda(inputs, self.qweight, self.scales, self.qzeros, self.group_size,
                                                         self.split_k_iters, w_bit=self.w_bit)
        else:
            out = awq_inference_engine.gemv_forward_cuda(inputs, self.qweight, self.scales, self.qzeros, self.w_bit,
                                                         self.group_size)
        out = out + self.bias if self.bias is not None else out
        return out.reshape(out_shape)

    def extra_repr(self) -> str:
        return 'in_features={}, out_features={}, bias={}, w_bit={}, group_size={}'.format(
            self.in_features, self.out_features, self.bias is not None, self.w_bit, self.group_size
        )
