"""Drop-in for the reference's pybind extension `awq_inference_engine` (inference/kernels/csrc/pybind.cpp:9-18),
restricted to the quantisation entry points on the hot path: gemv_forward_cuda (gemv_cuda.h:4-10) and
gemm_forward_cuda (gemm_cuda.h:3-4).  Both land in libbdq.so; outputs are allocated here (the reference allocates
them inside the extension, gemv_cuda.cu:305-308)."""
import torch

from . import _lib as L


def _check_fp16(*ts):
    for t in ts:
        if t.dtype != torch.float16:
            raise TypeError(f"expected a float16 tensor (the reference kernels are fp16-only), got {t.dtype}")


def gemv_forward_cuda(in_feats, kernel, scaling_factors, zeros, w_bit, group_size, pdl=False):
    """[B, IC] fp16 x packed [OC, IC/16] int32 -> [B, OC] fp16.  Unlike the reference, unsupported (w_bit, group_size)
    raise instead of returning uninitialised memory (gemv_cuda.cu:316-344)."""
    L.require_cuda(in_feats, kernel, scaling_factors, zeros)
    _check_fp16(in_feats, scaling_factors)
    if w_bit != 2:
        raise NotImplementedError("bitdistiller_b200 implements the 2-bit packed GEMV (the 4-bit kernels of the reference "
                                  "are not a BASELINE config)")
    x = in_feats.contiguous()
    B, IC = x.shape
    OC = kernel.shape[0]
    assert kernel.is_contiguous() and scaling_factors.is_contiguous() and zeros.is_contiguous()
    assert kernel.shape[1] * 16 == IC, "qweight must be [OC, IC/16] int32"
    out = torch.empty((B, OC), dtype=torch.float16, device=x.device)
    with torch.cuda.device(x.device):
        rc = L.lib().bdq_gemv_w2(L.ptr(x), L.ptr(kernel), L.ptr(scaling_factors), L.ptr(zeros), L.ptr(out), L.i64(B),
                                 L.i64(IC), L.i64(OC), int(group_size), 1 if pdl else 0, L.stream_ptr(x.device))
    L.check(rc, "bdq_gemv_w2")
    return out


def gemm_forward_cuda(in_feats, kernel, scaling_factors, zeros, group_size, split_k_iters, w_bit=2):
    """[M, IC] fp16 x packed 2-bit weights -> [M, OC] fp16 on the tcgen05 GEMM (the reference's CUDA GEMM is 4-bit only,
    gemm_cuda_gen.cu:31-32; its 2-bit prefill path is the Triton kernel this replaces)."""
    from .triton_kernels import gemm_w2
    if w_bit != 2:
        raise NotImplementedError("only the 2-bit packed GEMM is implemented")
    return gemm_w2(in_feats, kernel, scaling_factors, zeros, group_size)
