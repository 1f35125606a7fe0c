"""GPU parity tests for the packed int2 prefill GEMM (tcgen05 path), through the C ABI.
Bar (BASELINE.json north_star): GEMM outputs within rtol 1e-2 of the reference math (dequant rounded to fp16, fp32
accumulate, fp16 store = quant_matmul_kernel, triton_kernels.py:89-93)."""
import numpy as np
import pytest
import torch

import oracle
from _util import make_weight

pytestmark = pytest.mark.gpu
dev = torch.device("cuda:0")


def _packed_layer(oc, ic, seed, G=128):
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq
    w = make_weight("randn", oc, ic, torch.float16, seed).to(dev)
    dq, s, z, _ = Q.int_fakequant(w, 2, G, 0, want_scale_zp=True)
    return pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, G), dq


def _ref_gpu(x, qw, qz, qs, G):
    """dequant rounded to fp16 then fp32 matmul, on the GPU (cuBLAS fp32) -- the full-size check."""
    oc, wpr = qw.shape
    ic = wpr * 16
    sh = (torch.arange(16, device=dev, dtype=torch.int32) * 2)
    codes = ((qw[:, :, None] >> sh) & 3).reshape(oc, ic).float()
    ng = ic // G
    zz = ((qz[:, :, None] >> sh) & 3).reshape(oc, -1)[:, :ng].float()
    dqw = (qs[:, :ng].float().repeat_interleave(G, 1) * (codes - zz.repeat_interleave(G, 1))).half()
    return (x.float() @ dqw.float().T), dqw


def _check(out, ref, x, dqw):
    out, ref = out.double(), ref.double()
    scale = (x.double().abs() @ dqw.double().abs().T)
    tol = 1e-2 * ref.abs() + 2e-3 * scale / np.sqrt(x.shape[1]) + 1e-6
    bad = (out - ref).abs() > tol
    assert not bool(bad.any()), (int(bad.sum()), float((out - ref).abs().max()))
    assert float((out - ref).norm() / ref.norm()) < 1e-3


@pytest.mark.parametrize("m,oc,ic,G", [(128, 256, 128, 128), (256, 512, 1024, 128), (2048, 4096, 4096, 128),
                                       (200, 256, 11008, 128), (77, 1024, 4096, 128), (384, 512, 2048, 64),
                                       (300, 128, 4096, 128), (130, 352, 1024, 128), (64, 64, 512, 128),   # partial last N tile
                                       # two token tiles per CTA (grids that cover >= half the SMs): ragged second tile
                                       (257, 12288, 512, 128), (129, 20480, 256, 128), (1000, 4096, 1024, 64)])
def test_gemm_w2_tcgen05_vs_reference_math(m, oc, ic, G):
    from bitdistiller_b200.triton_kernels import gemm_w2
    (qw, qz, qs), dq = _packed_layer(oc, ic, 5, G)
    x = torch.randn(m, ic, generator=torch.Generator().manual_seed(m)).to(torch.float16).to(dev)
    out = gemm_w2(x, qw, qs, qz, G)
    torch.cuda.synchronize()
    assert out.shape == (m, oc) and out.dtype == torch.float16
    ref, dqw = _ref_gpu(x, qw, qz, qs, G)
    _check(out, ref, x, dqw)
    # a slab against the CPU oracle (double accumulation of the fp16-rounded dequant)
    ms, os_ = min(m, 16), min(oc, 64)
    orc = oracle.wq_matmul(x[:ms].cpu(), qw[:os_].cpu(), qz[:os_].cpu(), qs[:os_].cpu(), G, 2, dequant_fp16=True)
    _check(out[:ms, :os_].cpu(), orc, x[:ms].cpu(), dqw[:os_].cpu())


def test_wqlinear_prefill_and_quant_matmul_v2_convention():
    from bitdistiller_b200.qmodule import WQLinear
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    from bitdistiller_b200.triton_kernels import quant_matmul_v2
    lin = torch.nn.Linear(1024, 512, bias=True, dtype=torch.float16, device=dev)
    lin.weight.data = make_weight("randn", 512, 1024, torch.float16, 9).to(dev)
    lin.weight.data, s, z = pseudo_quantize_tensor(lin.weight.data, n_bit=2, q_group_size=128, get_scale_zp=True)
    ql = WQLinear.from_linear(lin, 2, 128, False, s, z)
    x = torch.randn(2, 40, 1024, device=dev, dtype=torch.float16)
    with torch.no_grad():
        y = ql(x)
        ref = torch.nn.functional.linear(x.float(), lin.weight.data.float(), lin.bias.float())
    assert y.shape == (2, 40, 512)
    assert float((y.float() - ref).norm() / ref.norm()) < 2e-3
    # the reference's transposed calling convention (qmodule.py:154-166)
    y2 = quant_matmul_v2(x.reshape(-1, 1024), ql.qweight.T.contiguous(), ql.qzeros.T.contiguous(), ql.scales.T.contiguous(),
                         M=80, N=512, K=1024, pack_num=16, group_size=128, w_bit=2, offset=3)
    assert torch.equal(y2 + ql.bias, y.reshape(-1, 512))


def test_gemm_w2_fallback_shapes():
    """OC not a multiple of 32 takes the GEMV row-block fallback and still matches."""
    from bitdistiller_b200.triton_kernels import gemm_w2
    (qw, qz, qs), dq = _packed_layer(48, 512, 3)
    x = torch.randn(20, 512, generator=torch.Generator().manual_seed(1)).to(torch.float16).to(dev)
    out = gemm_w2(x, qw, qs, qz, 128)
    ref, dqw = _ref_gpu(x, qw, qz, qs, 128)
    _check(out, ref, x, dqw)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_second_device_in_one_process():
    """MaxDynamicSharedMemorySize is a per-device attribute -- the > 48 KB kernels (tcgen05 GEMM, stream GEMV,
    autoclip) must also launch on cuda:1 after they ran on cuda:0 in the same process."""
    from bitdistiller_b200 import awq_inference_engine as E
    from bitdistiller_b200.autoclip import auto_2clip_layer
    from bitdistiller_b200.triton_kernels import gemm_w2
    (qw, qz, qs), dq = _packed_layer(512, 1024, 5)
    x = torch.randn(256, 1024, generator=torch.Generator().manual_seed(1)).to(torch.float16)
    w = make_weight("randn", 64, 256, torch.bfloat16, 3)
    xa = torch.randn(512, 256, generator=torch.Generator().manual_seed(2)).to(torch.bfloat16)
    outs = []
    for d in ("cuda:0", "cuda:1"):
        o = gemm_w2(x.to(d), qw.to(d), qs.to(d), qz.to(d), 128)
        v = E.gemv_forward_cuda(x[:1].to(d), qw.to(d), qs.to(d), qz.to(d), 2, 128)
        a, b = auto_2clip_layer(w.to(d), xa.to(d), 2, {"zero_point": True, "q_group_size": 128, "quant_type": "int"})
        torch.cuda.synchronize(d)
        outs.append((o.cpu(), v.cpu(), a.cpu(), b.cpu()))
    for p, q in zip(*outs):
        assert torch.equal(p, q)


@pytest.mark.parametrize("m,oc,ic", [(128, 256, 64), (300, 1024, 4096), (2048, 4096, 4096), (77, 352, 1024), (257, 12288, 512), (9, 32, 128)])
def test_gemm_f16_tcgen05_vs_fp32_matmul(m, oc, ic):
    """bdq_gemm_f16: the same tcgen05 kernel with the B tile loaded by TMA from fp16 weights [OC, IC] (WQLinearNF3's prefill):
    within rtol 1e-2 of the fp32 matmul of the same fp16 operands (partial last N tile, ragged M, both token-tile counts)."""
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200.nf3_packed import gemm_f16
    g = torch.Generator().manual_seed(m + oc)
    w = (torch.randn(oc, ic, generator=g) * 0.02).half().to(dev)
    x = torch.randn(m, ic, generator=g).half().to(dev)
    n0 = L.launch_count()
    out = gemm_f16(x, w)
    torch.cuda.synchronize()
    assert L.launch_count() == n0 + 1 and out.shape == (m, oc) and out.dtype == torch.float16
    ref = x.float() @ w.float().T
    _check(out, ref, x, w)
    with pytest.raises(L.BdqError):
        gemm_f16(x[:, :ic - 8].contiguous(), w[:, :ic - 8].contiguous())


@pytest.mark.parametrize("tile", ["auto", "11", "21", "12", "22"])
def test_gemm_partial_tiles_write_nothing_outside(tile, monkeypatch):
    """Partial last tiles in both directions (M = 200: a 128-row tile + 72 rows; OC = 1056 = 8 x 128 + 32) through the C ABI with
    the output inside a larger sentinel-filled buffer: the values match an fp32 matmul of the dequantised weights and not one
    element around the output is touched.  BDQ_GEMM_TILE is read once per process, so the forced tiles run in a subprocess."""
    import subprocess
    import sys
    code = r"""
import ctypes, torch
from bitdistiller_b200 import _lib as L, quantizer as Q
from bitdistiller_b200.qmodule import pack_wq
dev = torch.device("cuda:0")
M, IC, OC, PAD = 200, 1024, 1056, 4096
g = torch.Generator(device=dev).manual_seed(5)
w = (torch.randn(OC, IC, device=dev, generator=g) * 0.03).half()
dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
qw, qz, qs = pack_wq(dq, s.view(OC, -1), z.view(OC, -1), 2, 128)
x = torch.randn(M, IC, device=dev, generator=g).half()
big = torch.full((PAD + M * OC + PAD,), 777.0, dtype=torch.float16, device=dev)
out = big[PAD:PAD + M * OC]
rc = L.lib().bdq_gemm_w2(L.ptr(x), L.ptr(qw), L.ptr(qs), L.ptr(qz), ctypes.c_void_p(out.data_ptr()), L.i64(M), L.i64(IC), L.i64(OC), 128,
                         L.stream_ptr(dev))
L.check(rc, "bdq_gemm_w2")
torch.cuda.synchronize()
ref = x.float() @ dq.float().T
err = float((out.view(M, OC).float() - ref).norm() / ref.norm())
assert err < 2e-3, err
assert bool((big[:PAD] == 777.0).all()) and bool((big[PAD + M * OC:] == 777.0).all()), "wrote outside the output"
print("ok", err)
"""
    import os
    env = dict(os.environ)
    env.pop("BDQ_GEMM_TILE", None)
    if tile != "auto":
        env["BDQ_GEMM_TILE"] = tile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], cwd=root, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-500:] + r.stderr[-1500:]
