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


GEMV_PDL, GEMV_CTAS1, GEMV_RING, GEMV_SILU_MUL = 1, 2, 4, 64   # include/bdq.h


def gemv_forward_cuda(in_feats, kernel, scaling_factors, zeros, w_bit, group_size, pdl=False, flags=0):
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
    with L.on(x.device):
        rc = L.lib().bdq_gemv_w2(L.ptr(x), L.ptr(kernel), L.ptr(scaling_factors), L.ptr(zeros), L.ptr(out), L.i64(B),
                                 L.i64(IC), L.i64(OC), int(group_size), int(flags) | (GEMV_PDL if pdl else 0),
                                 L.stream_ptr(x.device))
    L.check(rc, "bdq_gemv_w2")
    return out


def gemv_multi_forward_cuda(in_feats, kernels, scaling_factors, zeros, group_size, pdl=False, flags=0, outs=None,
                            prefetch=()):
    """Batch-1 decode GEMV of up to 4 packed matrices sharing `in_feats` [1, IC] in ONE launch (bdq_gemv_w2_multi): the
    fused q/k/v and gate/up projections of inference/modules/fused_attn.py:304-340 / fused_mlp.py:44-114 without
    concatenating the packed tensors.  `prefetch`: up to 4 tensors (the packed weights of the next launch of the decode
    chain) to pull into L2 while this launch computes.  Returns a list of [1, OC_i] fp16 tensors."""
    import ctypes
    n = len(kernels)
    L.require_cuda(in_feats, *kernels, *scaling_factors, *zeros)
    _check_fp16(in_feats, *scaling_factors)
    x = in_feats.contiguous()
    if x.numel() != x.shape[-1]:
        raise ValueError("gemv_multi_forward_cuda is the batch-1 decode path: in_feats must hold one row")
    IC = x.shape[-1]
    for k, s, z in zip(kernels, scaling_factors, zeros):
        assert k.is_contiguous() and s.is_contiguous() and z.is_contiguous() and k.shape[1] * 16 == IC
    if outs is None:
        outs = [torch.empty((1, k.shape[0]), dtype=torch.float16, device=x.device) for k in kernels]
    arr = lambda ts: (ctypes.c_void_p * n)(*[t.data_ptr() for t in ts])
    oc = (ctypes.c_int64 * n)(*[k.shape[0] for k in kernels])
    npf = len(prefetch)
    pf_ptr = (ctypes.c_void_p * max(npf, 1))(*[t.data_ptr() for t in prefetch])
    pf_len = (ctypes.c_int64 * max(npf, 1))(*[t.numel() * t.element_size() for t in prefetch])
    with L.on(x.device):
        rc = L.lib().bdq_gemv_w2_multi(L.ptr(x), n, arr(kernels), arr(scaling_factors), arr(zeros), arr(outs), oc,
                                       L.i64(IC), int(group_size), int(flags) | (GEMV_PDL if pdl else 0),
                                       npf, pf_ptr, pf_len, L.stream_ptr(x.device))
    L.check(rc, "bdq_gemv_w2_multi")
    return outs


def gemv_silu_mul_forward_cuda(in_feats, gate, up, group_size, pdl=False, flags=0, out=None):
    """Batch-1 fused MLP front half (inference/modules/fused_mlp.py:44-108, rows == 1): silu(x @ gate^T) * (x @ up^T) in
    ONE launch.  `gate` / `up` = (qweight, scales, qzeros) of two packed matrices with equal out_features; the kernel
    walks their 16-channel blocks interleaved and applies the reference's three fp16 roundings (after the GEMV, after
    F.silu, after the product) in the split-K epilogue.  Returns [1, OC] fp16."""
    import ctypes
    L.require_cuda(in_feats, *gate, *up)
    _check_fp16(in_feats, gate[1], up[1])
    x = in_feats.contiguous()
    if x.numel() != x.shape[-1]:
        raise ValueError("gemv_silu_mul_forward_cuda is the batch-1 decode path: in_feats must hold one row")
    IC, OC = x.shape[-1], gate[0].shape[0]
    if up[0].shape[0] != OC:
        raise ValueError("gate and up must have the same out_features")
    for k, s, z in (gate, up):
        assert k.is_contiguous() and s.is_contiguous() and z.is_contiguous() and k.shape[1] * 16 == IC
    if out is None:
        out = torch.empty((1, OC), dtype=torch.float16, device=x.device)
    arr = lambda a, b: (ctypes.c_void_p * 2)(a.data_ptr(), b.data_ptr())
    outs = (ctypes.c_void_p * 2)(out.data_ptr(), None)
    oc = (ctypes.c_int64 * 2)(OC, OC)
    with L.on(x.device):
        rc = L.lib().bdq_gemv_w2_multi(L.ptr(x), 2, arr(gate[0], up[0]), arr(gate[1], up[1]), arr(gate[2], up[2]), outs, oc,
                                       L.i64(IC), int(group_size), int(flags) | GEMV_SILU_MUL | (GEMV_PDL if pdl else 0),
                                       0, None, None, L.stream_ptr(x.device))
    L.check(rc, "bdq_gemv_w2_multi(silu_mul)")
    return out


def gemm_forward_cuda(in_feats, kernel, scaling_factors, zeros, group_size, split_k_iters, w_bit=2):
    """[M, IC] fp16 x packed 2-bit weights -> [M, OC] fp16 on the tcgen05 GEMM (the reference's CUDA GEMM is 4-bit only,
    gemm_cuda_gen.cu:31-32; its 2-bit prefill path is the Triton kernel this replaces)."""
    from .triton_kernels import gemm_w2
    if w_bit != 2:
        raise NotImplementedError("only the 2-bit packed GEMM is implemented")
    return gemm_w2(in_feats, kernel, scaling_factors, zeros, group_size)


class _Phase(__import__("ctypes").Structure):
    """bdq_gemv_phase_t (include/bdq.h)."""
    import ctypes as _c
    _fields_ = [("x", _c.c_void_p), ("nmat", _c.c_int), ("group", _c.c_int), ("flags", _c.c_int), ("reserved", _c.c_int),
                ("IC", _c.c_int64), ("qweight", _c.c_void_p * 4), ("scales", _c.c_void_p * 4), ("qzeros", _c.c_void_p * 4),
                ("out", _c.c_void_p * 4), ("OC", _c.c_int64 * 4)]


GEMV_CHAIN_NOWAIT = 256


class DecodeChain:
    """A decoded token's batch-1 GEMVs as ONE persistent launch (bdq_gemv_w2_chain_*): the per-linear
    `gemv_forward_cuda` calls of WQLinear.forward (quantization/qmodule.py:166) for a whole token, grouped into phases of
    up to 4 matrices that share their activation vector (fused q/k/v, gate/up: inference/modules/fused_attn.py:304-340,
    fused_mlp.py:44-114).  `phases`: list of dicts
        {"x": [1, IC] fp16 tensor, "mats": [(qweight, scales, qzeros, out [1, OC] fp16), ...], "silu": False, "nowait": False}
    executed in order; phase p reads x after every earlier phase has stored its outputs, so x may be an `out` of an
    earlier phase ("nowait": x is an external input, no dependency).  run() is bit-equal to issuing the phases as
    gemv_multi_forward_cuda / gemv_silu_mul_forward_cuda launches one after the other.  The tensors are referenced by
    address: keep them alive and do not reallocate them."""

    def __init__(self, phases, group_size=128, max_phases_per_launch=None):
        import ctypes
        if not phases:
            raise ValueError("DecodeChain needs at least one phase")
        self._keep, self._runs = [], []
        dev = phases[0]["x"].device
        for ph in phases:
            x, mats = ph["x"], ph["mats"]
            L.require_cuda(x, *[t for m in mats for t in m if t is not None])
            _check_fp16(x, *[m[1] for m in mats])
            if x.numel() != x.shape[-1] or not x.is_contiguous():
                raise ValueError("DecodeChain is the batch-1 decode path: every x must hold one contiguous row")
            if not 1 <= len(mats) <= 4:
                raise ValueError("a phase takes 1..4 matrices")
        self.n, self.device = len(phases), dev
        self.max_ic = max(ph["x"].shape[-1] for ph in phases)
        lib = L.lib()
        lib.bdq_gemv_w2_chain_workspace_bytes.restype = ctypes.c_int64
        cap = int(lib.bdq_gemv_w2_chain_max_phases(L.i64(self.max_ic)))
        if cap < 1:
            raise L.BdqError(f"DecodeChain: in_features {self.max_ic} do not fit the persistent kernel's shared memory")
        if max_phases_per_launch:
            cap = max(1, min(cap, int(max_phases_per_launch)))
        # one launch holds `cap` phases (the plan lives in shared memory): longer chains run as consecutive launches
        for s0 in range(0, self.n, cap):
            sub = phases[s0:s0 + cap]
            arr = (_Phase * len(sub))()
            for a, ph in zip(arr, sub):
                x, mats = ph["x"], ph["mats"]
                a.x, a.nmat, a.group, a.IC = x.data_ptr(), len(mats), int(group_size), x.shape[-1]
                a.flags = (GEMV_SILU_MUL if ph.get("silu") else 0) | (GEMV_CHAIN_NOWAIT if ph.get("nowait") else 0)
                for i, (qw, sc, qz, out) in enumerate(mats):
                    assert qw.is_contiguous() and sc.is_contiguous() and qz.is_contiguous() and qw.shape[1] * 16 == x.shape[-1]
                    a.qweight[i], a.scales[i], a.qzeros[i] = qw.data_ptr(), sc.data_ptr(), qz.data_ptr()
                    a.out[i] = 0 if out is None else out.data_ptr()
                    a.OC[i] = qw.shape[0]
                    self._keep += [qw, sc, qz, out]
                self._keep.append(x)
            nbytes = int(lib.bdq_gemv_w2_chain_workspace_bytes(len(sub)))
            ws = torch.zeros(nbytes + 256, dtype=torch.uint8, device=dev)
            ptr = ctypes.c_void_p(ws.data_ptr() + (-ws.data_ptr()) % 256)
            with L.on(dev):
                rc = lib.bdq_gemv_w2_chain_prepare(arr, len(sub), ptr, L.i64(nbytes), L.stream_ptr(dev))
            L.check(rc, "bdq_gemv_w2_chain_prepare")
            self._runs.append((ws, ptr, len(sub)))

    def run(self, lookahead=0):
        with L.on(self.device):
            for _, ptr, n in self._runs:
                rc = L.lib().bdq_gemv_w2_chain_run(ptr, n, L.i64(self.max_ic), int(lookahead), L.stream_ptr(self.device))
                L.check(rc, "bdq_gemv_w2_chain_run")
