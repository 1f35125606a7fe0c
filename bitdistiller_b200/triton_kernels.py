"""Drop-in for the entry points of the reference's quantization/triton_kernels.py -- with no Triton behind them.

quant_matmul_v2 (triton_kernels.py:102-134) is the 2-bit prefill GEMM of the reference; here it lands in bdq_gemm_w2
(tcgen05 tensor cores, TMA-fed activations, in-register dequant).  The reference passes TRANSPOSED contiguous copies of
the packed tensors (qmodule.py:154-158); that calling convention is accepted for compatibility (the operands are
transposed back, which costs what the reference pays on every call), while WQLinear.forward of this package calls gemm_w2
with the row-major buffers directly and copies nothing.  quant_gemv_v2 / quant_gemv_v3 are dead code in the reference
(never called); the names are kept and route to the packed GEMV."""
import torch

from . import _lib as L


def gemm_w2(x, qweight, scales, qzeros, group_size):
    """x [M, IC] fp16, qweight int32 [OC, IC/16], scales fp16 [OC, zw*16], qzeros int32 [OC, zw] -> [M, OC] fp16."""
    L.require_cuda(x, qweight, scales, qzeros)
    if x.dtype != torch.float16 or scales.dtype != torch.float16:
        raise TypeError("gemm_w2 expects float16 activations and scales (the reference kernels are fp16-only)")
    x = x.contiguous()
    M, IC = x.shape
    OC = qweight.shape[0]
    assert qweight.is_contiguous() and scales.is_contiguous() and qzeros.is_contiguous()
    assert qweight.shape[1] * 16 == IC, "qweight must be [OC, IC/16] int32"
    out = torch.empty((M, OC), dtype=torch.float16, device=x.device)
    with L.on(x.device):
        rc = L.lib().bdq_gemm_w2(L.ptr(x), L.ptr(qweight), L.ptr(scales), L.ptr(qzeros), L.ptr(out), L.i64(M), L.i64(IC),
                                 L.i64(OC), int(group_size), L.stream_ptr(x.device))
    L.check(rc, "bdq_gemm_w2")
    return out


def quant_matmul_v2(a, qw, qzeros, scales, *, M, N, K, pack_num, group_size, w_bit, offset):
    """triton_kernels.py:102-134 calling convention: qw [K/pack, N], qzeros [ceil(K/G/pack), N], scales [K/G (padded), N]."""
    assert w_bit == 2 and pack_num == 16, "only the 2-bit packed format is implemented"
    assert a.shape[-1] == K and qw.shape == (K // pack_num, N)
    return gemm_w2(a.reshape(-1, K), qw.T.contiguous(), scales.T.contiguous(), qzeros.T.contiguous(), group_size)


def quant_gemv_v2(a, qw, qzeros, scales, *, M, N, K, pack_num, group_size, w_bit, offset):
    from .awq_inference_engine import gemv_forward_cuda
    assert w_bit == 2 and pack_num == 16
    return gemv_forward_cuda(a.reshape(-1, K), qw.T.contiguous(), scales.T.contiguous(), qzeros.T.contiguous(), 2, group_size)


quant_gemv_v3 = quant_gemv_v2
