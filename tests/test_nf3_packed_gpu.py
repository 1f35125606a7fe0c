"""Packed NF3 (this build's own format: the reference has none, SURVEY section 8c -- parity is pinned on VALUES):
dequant(pack(w)) must equal the reference NF3 fake-quantiser's output bit for bit in fp16, on the committed golden
fake-quant cases and on large random tensors; the decode GEMV is checked against a float64 matmul of those values."""
import numpy as np
import pytest
import torch

import oracle
from _util import bits_equal, load_cases, make_weight, t

pytestmark = pytest.mark.gpu
dev = torch.device("cuda:0")


def _unpack_reference(qweight, qhigh):
    """numpy restatement of the layout in include/bdq.h -> level index e [OC, IC]."""
    qw = qweight.cpu().numpy().astype(np.uint32)
    qh = qhigh.cpu().numpy().astype(np.uint32)
    oc, ic = qw.shape[0], qw.shape[1] * 16
    k = np.arange(ic)
    lo = (qw[:, k // 16] >> (2 * (k % 16)).astype(np.uint32)) & 3
    kk = k % 32   # nf3_packed.cuh: channel 16w + 2n + p of a 32-channel chunk keeps its high bit at 4n + 2w + p
    pos = 4 * ((kk % 16) // 2) + 2 * (kk // 16) + (kk % 2)
    hi = (qh[:, k // 32] >> pos.astype(np.uint32)) & 1
    return (lo | (hi << 2)).astype(np.int64)


@pytest.mark.parametrize("kind", ["randn", "uniform", "heavy"])
@pytest.mark.parametrize("G", [128, 64, 32])
def test_pack_dequant_equals_fake_quant(kind, G):
    from bitdistiller_b200 import nf3_packed as P
    from bitdistiller_b200.quantizer import pseudo_quantize_n2f3_tensor
    w = make_weight(kind, 384, 1024, torch.float16, 77).to(dev)
    qw, qh, sc = P.pack_nf3(w, G)
    assert qw.shape == (384, 64) and qh.shape == (384, 32) and sc.shape == (384, 2 * 1024 // G)
    dq = P.dequant_nf3(qw, qh, sc, G)
    ref = oracle.nf3_fakequant(w.cpu(), G)                       # CPU oracle of quantizer.py:184-242
    assert bits_equal(dq.cpu(), ref["dq"])
    assert bits_equal(dq, pseudo_quantize_n2f3_tensor(w, G))   # and the drop-in fake-quantiser itself
    # the layout: level index e -> the level table, times the sign's scale
    e = _unpack_reference(qw, qh)
    lv = torch.tensor([-1.0, -0.5350227355957031, -0.2469314038753510, 0.0, 0.1833375245332718, 0.3819939494132996,
                       0.6229856610298157, 1.0]).half().float().numpy()
    s = sc.float().cpu().numpy().reshape(384, -1, 2)
    scale = np.where(e < 3, np.repeat(s[:, :, 1], G, axis=1), np.repeat(s[:, :, 0], G, axis=1))
    manual = torch.from_numpy((lv[e] * scale).astype(np.float32)).half()
    assert bits_equal(dq.cpu(), manual)


def test_pack_dequant_golden_cases():
    """Every fp16 NF3 case of the committed golden fake-quant vectors (generated from the reference's own code)."""
    from bitdistiller_b200 import nf3_packed as P
    n = 0
    for meta, c in load_cases("ste"):
        qname, dname, G = meta[0], meta[1], int(meta[3])
        if qname != "nf3" or dname != "f16" or G % 32:
            continue
        w = t(c["w"], torch.float16).to(dev)
        if w.shape[-1] % G or w.shape[-1] % 32:
            continue
        w2 = w.reshape(-1, w.shape[-1])
        dq = P.dequant_nf3(*P.pack_nf3(w2, G), G)
        assert bits_equal(dq.cpu().reshape(w.shape), t(c["y"], torch.float16)), meta
        n += 1
    assert n >= 3


@pytest.mark.parametrize("B", [1, 3, 8])
@pytest.mark.parametrize("oc,ic", [(1024, 4096), (48, 2176), (16, 128)])
def test_gemv_nf3(B, oc, ic):
    from bitdistiller_b200 import nf3_packed as P
    w = (make_weight("randn", oc, ic, torch.float16, 5).float() * (ic ** -0.5 / 0.02)).half().to(dev)
    qw, qh, sc = P.pack_nf3(w, 128)
    dq = P.dequant_nf3(qw, qh, sc, 128)
    x = torch.randn(B, ic, generator=torch.Generator().manual_seed(3)).half().to(dev)
    y = P.gemv_nf3(x, qw, qh, sc, 128)
    ref = x.double() @ dq.double().T
    assert y.shape == (B, oc) and y.dtype == torch.float16
    scale = x.double().abs() @ dq.double().abs().T
    tol = 1e-2 * ref.abs() + 2e-3 * scale / np.sqrt(ic) + 1e-6
    assert not bool(((y.double() - ref).abs() > tol).any())
    assert float((y.double() - ref).norm() / ref.norm()) < 2e-3


def test_wqlinear_nf3_module():
    from bitdistiller_b200.nf3_packed import WQLinearNF3
    from bitdistiller_b200.quantizer import pseudo_quantize_n2f3_tensor
    lin = torch.nn.Linear(1024, 512, bias=True, dtype=torch.float16, device=dev)
    lin.weight.data = (make_weight("randn", 512, 1024, torch.float16, 9).float() * (1024 ** -0.5 / 0.02)).half().to(dev)
    m = WQLinearNF3.from_linear(lin, 128)
    assert set(m.state_dict()) == {"qweight", "qhigh", "scales", "bias"}
    dq = pseudo_quantize_n2f3_tensor(lin.weight.data, 128).float()
    for rows in (1, 6, 40):
        x = torch.randn(2, rows, 1024, device=dev, dtype=torch.float16)
        y = m(x)
        ref = torch.nn.functional.linear(x.float(), dq, lin.bias.float())
        assert y.shape == (2, rows, 512)
        assert float((y.float() - ref).norm() / ref.norm()) < 3e-3
    skel = WQLinearNF3.from_linear(lin, 128, init_only=True)
    skel.load_state_dict(m.state_dict())
    assert torch.equal(skel(x), m(x))
    with pytest.raises(TypeError):
        from bitdistiller_b200.nf3_packed import pack_nf3
        pack_nf3(lin.weight.data.float(), 128)


@pytest.mark.parametrize("oc,ic,kind", [(4096, 4096, "randn"), (512, 12288, "heavy"), (2048, 11008, "uniform"), (16, 512, "randn"),
                                        (32000, 4096, "randn")])
def test_gemv_nf3_stream_kernel(oc, ic, kind):
    """The PRMT-lookup / INT8-MMA streaming kernel (group 128, OC % 16 == 0) on the decode shapes: skewed groups (sp != sn, both
    signs of the level table and both scales in use), an outlier activation channel (block fixed point of x), against a
    float64 matmul of the dequantised weights -- and against the pair-lookup kernel on the same tensors."""
    from bitdistiller_b200 import nf3_packed as P
    w = make_weight(kind, oc, ic, torch.float16, 11).float() * (ic ** -0.5 / 0.02)
    w = (w + 0.4 * w.std() * (torch.arange(oc).view(-1, 1) % 3 - 1)).half().to(dev)      # rows skewed down / centred / up
    qw, qh, sc = P.pack_nf3(w, 128)
    dq = P.dequant_nf3(qw, qh, sc, 128)
    x = torch.randn(1, ic, generator=torch.Generator().manual_seed(8)).half()
    x[0, 7] = 40.0
    x = x.to(dev)
    y = P.gemv_nf3(x, qw, qh, sc, 128)
    ref = x.double() @ dq.double().T
    scale = x.double().abs() @ dq.double().abs().T
    tol = 1e-2 * ref.abs() + 2e-3 * scale / np.sqrt(ic) + 1e-6
    assert not bool(((y.double() - ref).abs() > tol).any())
    assert float((y.double() - ref).norm() / ref.norm()) < 1e-3
    # the other kernel (group 64 re-pack is not the same tensor, so force it by an OC that is not a multiple of 16)
    y2 = P.gemv_nf3(x, qw[:oc - 8], qh[:oc - 8], sc[:oc - 8], 128)
    assert float((y2.double() - y[:, :oc - 8].double()).norm() / ref.norm()) < 1e-3


def test_gemv_nf3_stream_nonfinite():
    from bitdistiller_b200 import nf3_packed as P
    w = (make_weight("randn", 64, 1024, torch.float16, 3).float() * 1.5).half().to(dev)
    qw, qh, sc = P.pack_nf3(w, 128)
    x = torch.randn(1, 1024, generator=torch.Generator().manual_seed(1)).half().to(dev)
    sc2 = sc.clone()
    sc2[5, 6] = float("nan")                      # sp of group 3 of channel 5
    y = P.gemv_nf3(x, qw, qh, sc2, 128)
    assert bool(y[0, 5].isnan()) and int(y.isnan().sum()) == 1
    x2 = x.clone()
    x2[0, 300] = float("inf")
    assert bool(P.gemv_nf3(x2, qw, qh, sc, 128).isnan().all())
    z = torch.zeros_like(sc)                      # all-zero scales: zero output, not NaN
    assert float(P.gemv_nf3(x, qw, qh, z, 128).abs().max()) == 0.0


def test_gemv_nf3_writes_nothing_outside_the_output():
    """Both NF3 GEMV kernels through the C ABI with the output inside a sentinel-filled buffer (ragged OC for the pair lookup,
    a partial last K window for the streaming kernel)."""
    import ctypes

    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import nf3_packed as P
    for oc, ic in [(1040, 2176), (1038 // 2 * 2 - 2, 1024)]:     # 1040 = 65 x 16: streaming kernel; 1036: pair lookup
        w = (make_weight("randn", oc + (-oc) % 2, ic, torch.float16, 9).float() * 1.7).half().to(dev)[:oc].contiguous()
        qw, qh, sc = P.pack_nf3(w, 128)
        dq = P.dequant_nf3(qw, qh, sc, 128)
        B, PAD = 3, 512
        x = torch.randn(B, ic, generator=torch.Generator().manual_seed(2)).half().to(dev)
        big = torch.full((PAD + B * oc + PAD,), 777.0, dtype=torch.float16, device=dev)
        out = big[PAD:PAD + B * oc]
        rc = L.lib().bdq_gemv_nf3(L.ptr(x), L.ptr(qw), L.ptr(qh), L.ptr(sc), ctypes.c_void_p(out.data_ptr()), L.i64(B), L.i64(ic), L.i64(oc),
                                  128, L.stream_ptr(dev))
        L.check(rc, "bdq_gemv_nf3")
        torch.cuda.synchronize()
        ref = x.double() @ dq.double().T
        assert float((out.view(B, oc).double() - ref).norm() / ref.norm()) < 2e-3
        assert bool((big[:PAD] == 777.0).all()) and bool((big[PAD + B * oc:] == 777.0).all()), (oc, ic)
