"""GPU parity tests for the deploy side of the path: real-quantise + pack (WQLinear.from_linear) and the packed int2
GEMV, through the C ABI via the reference-named Python modules.  Packed bytes / zeros / scales: bit-exact.
GEMV outputs: rtol 1e-2 (BASELINE.json north_star), checked against the double-precision oracle."""
import numpy as np
import pytest
import torch

import oracle
from _util import DT, load_cases, make_weight, t

pytestmark = pytest.mark.gpu
dev = torch.device("cuda:0")


def test_pack_golden_bit_exact():
    from bitdistiller_b200.qmodule import WQLinear
    n = 0
    for meta, c in load_cases("pack"):
        dname, w_bit, G = meta[0], int(meta[1]), int(meta[2])
        dt = DT[dname]
        dq, s, z = t(c["dq"], dt).to(dev), t(c["s"], dt).to(dev), t(c["z"], dt).to(dev)
        rows, cols = dq.shape
        lin = torch.nn.Linear(cols, rows, bias=(n % 2 == 0), dtype=dt, device=dev)
        lin.weight.data = dq
        ql = WQLinear.from_linear(lin, w_bit, G, False, s, z)
        assert ql.qweight.dtype == torch.int32 and ql.qzeros.dtype == torch.int32 and ql.scales.dtype == torch.float16
        assert np.array_equal(ql.qweight.cpu().numpy(), c["qweight"]), meta
        assert np.array_equal(ql.qzeros.cpu().numpy(), c["qzeros"]), meta
        assert np.array_equal(ql.scales.float().cpu().numpy(), c["qscales"]), meta
        if lin.bias is not None:
            assert ql.bias.dtype == torch.float16 and torch.equal(ql.bias, lin.bias.half())
        n += 1
    assert n == 18


@pytest.mark.parametrize("dname", ["f16", "bf16", "f32"])
def test_pack_large_vs_oracle_and_roundtrip(dname):
    """[256, 11008] (ragged: 86 groups -> zeros width 6, scales padded to 96): bytes equal the oracle's, unpacked codes
    equal the fake-quant codes (format spec)."""
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq
    dt = DT[dname]
    w = make_weight("randn", 256, 11008, dt, 11)
    wd = w.to(dev)
    dq, s, z, codes = Q.int_fakequant(wd, 2, 128, 0, want_scale_zp=True, want_codes=True)
    qw, qz, qs = pack_wq(dq, s.view(256, -1), z.view(256, -1), 2, 128)
    rw, rz, rs = oracle.pack(dq.cpu(), s.view(256, -1).cpu(), z.view(256, -1).cpu(), 128, 2)
    assert torch.equal(qw.cpu(), rw) and torch.equal(qz.cpu(), rz) and torch.equal(qs.cpu(), rs)
    assert qz.shape == (256, 6) and qs.shape == (256, 96)
    un = (qw.cpu().numpy().view(np.uint32)[:, :, None] >> (np.arange(16, dtype=np.uint32) * 2)) & 3
    assert np.array_equal(un.reshape(256, 11008), codes.cpu().numpy())


def _packed_layer(oc, ic, seed, G=128):
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq
    w = make_weight("randn", oc, ic, torch.float16, seed).to(dev)
    dq, s, z, _ = Q.int_fakequant(w, 2, G, 0, want_scale_zp=True)
    return pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, G), dq


def _check_gemv(out, ref, x, dq):
    """rtol 1e-2 against the oracle, with an absolute floor tied to the fp16 output rounding and the size of the
    terms of the dot product (outputs near zero come from cancellation)."""
    out, ref = out.double().cpu(), ref.double().cpu()
    scale = (x.double().cpu().abs() @ dq.double().cpu().abs().T)  # sum_k |x_k w_k|
    tol = 1e-2 * ref.abs() + 2e-3 * scale / np.sqrt(x.shape[1]) + 1e-6
    bad = (out - ref).abs() > tol
    assert not bool(bad.any()), (int(bad.sum()), float((out - ref).abs().max()))
    # and tight in aggregate: relative Frobenius error far below the bar
    assert float((out - ref).norm() / ref.norm()) < 2e-3


@pytest.mark.parametrize("oc,ic", [(4096, 4096), (1024, 4096), (12288, 4096), (4096, 12288), (4096, 11008), (64, 128)])
@pytest.mark.parametrize("pdl", [False, True])
def test_gemv_w2_vs_oracle(oc, ic, pdl):
    from bitdistiller_b200 import awq_inference_engine as E
    (qw, qz, qs), dq = _packed_layer(oc, ic, 21)
    g = torch.Generator().manual_seed(5)
    x = torch.randn(1, ic, generator=g).to(torch.float16).to(dev)
    out = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128, pdl=pdl)
    assert out.shape == (1, oc) and out.dtype == torch.float16
    sub = slice(0, min(oc, 512))   # the CPU oracle is O(OC*IC): check a slab of channels plus the tail
    ref = oracle.wq_matmul(x.cpu(), qw[sub].cpu(), qz[sub].cpu(), qs[sub].cpu(), 128, 2)
    _check_gemv(out[:, sub], ref, x, dq[sub])
    ref_t = oracle.wq_matmul(x.cpu(), qw[-64:].cpu(), qz[-64:].cpu(), qs[-64:].cpu(), 128, 2)
    _check_gemv(out[:, -64:], ref_t, x, dq[-64:])
    # whole output against dequantised fp32 matmul on the GPU (same math, different summation order)
    full = (x.float() @ dq.float().T)
    assert float((out.float() - full).norm() / full.norm()) < 2e-3


@pytest.mark.parametrize("oc,ic,flags", [(4096, 4096, 2), (4096, 4096, 4), (4096, 12288, 6), (48, 2176, 0), (48, 2176, 2),
                                         (16, 128, 1), (1040, 640, 3), (4096, 4096, 16), (12288, 4096, 17), (4096, 12288, 17),
                                         (4096, 11008, 16), (48, 2176, 17), (16, 128, 16), (1040, 640, 18)])
def test_gemv_w2_batch1_variants(oc, ic, flags):
    """Batch 1 through both kernels: streaming (flags 0/1), one CTA per SM (2), the ring kernel (4), TMA weight tiles (16);
    ragged shapes
    (17 groups, 5 groups, a single group)."""
    from bitdistiller_b200 import awq_inference_engine as E
    (qw, qz, qs), dq = _packed_layer(oc, ic, 100 + flags)
    x = torch.randn(1, ic, generator=torch.Generator().manual_seed(oc + ic)).to(torch.float16).to(dev)
    out = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128, flags=flags)
    sub = slice(0, min(oc, 256))
    ref = oracle.wq_matmul(x.cpu(), qw[sub].cpu(), qz[sub].cpu(), qs[sub].cpu(), 128, 2)
    _check_gemv(out[:, sub], ref, x, dq[sub])
    full = (x.float() @ dq.float().T)
    assert float((out.float() - full).norm() / full.norm()) < 2e-3


@pytest.mark.parametrize("ocs,ic", [((4096, 1024, 1024), 4096), ((12288, 12288), 4096), ((4096, 48, 16, 1024), 11008)])
@pytest.mark.parametrize("pdl", [False, True])
@pytest.mark.parametrize("tma", [0, 16])
def test_gemv_w2_multi(ocs, ic, pdl, tma):
    """Fused q/k/v and gate/up launches: every matrix's output equals its stand-alone GEMV bit for bit (same kernel,
    same summation order) and the oracle within rtol 1e-2."""
    from bitdistiller_b200 import awq_inference_engine as E
    layers = [_packed_layer(oc, ic, 200 + i) for i, oc in enumerate(ocs)]
    x = torch.randn(1, ic, generator=torch.Generator().manual_seed(9)).to(torch.float16).to(dev)
    outs = E.gemv_multi_forward_cuda(x, [l[0][0] for l in layers], [l[0][2] for l in layers], [l[0][1] for l in layers],
                                     128, pdl=pdl, flags=tma, prefetch=[l[0][0] for l in layers[:2]] if pdl else ())
    for ((qw, qz, qs), dq), out in zip(layers, outs):
        single = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128)   # register-streaming variant: same arithmetic, same order
        assert torch.equal(out, single)
        sub = slice(0, min(qw.shape[0], 128))
        ref = oracle.wq_matmul(x.cpu(), qw[sub].cpu(), qz[sub].cpu(), qs[sub].cpu(), 128, 2)
        _check_gemv(out[:, sub], ref, x, dq[sub])


def test_gemv_w2_pdl_chain_dependent():
    """A dependent chain under programmatic dependent launch: each GEMV reads the previous one's output (x is loaded
    only after griddepcontrol.wait), results equal the plain-launch chain bit for bit."""
    from bitdistiller_b200 import awq_inference_engine as E
    (qw, qz, qs), _ = _packed_layer(4096, 4096, 77)
    x0 = torch.randn(1, 4096, generator=torch.Generator().manual_seed(3)).to(torch.float16).to(dev)

    def chain(pdl):
        x = x0
        for _ in range(6):
            x = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128, pdl=pdl)
            x = x  # [1, 4096] feeds the next launch
        return x
    a, b = chain(False), chain(True)
    assert torch.isfinite(a).all() and torch.equal(a, b)


@pytest.mark.parametrize("B", [2, 5, 8, 11])
def test_gemv_w2_batched(B):
    from bitdistiller_b200 import awq_inference_engine as E
    (qw, qz, qs), dq = _packed_layer(512, 4096, 33)
    x = torch.randn(B, 4096, generator=torch.Generator().manual_seed(B)).to(torch.float16).to(dev)
    out = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128)
    ref = oracle.wq_matmul(x.cpu(), qw.cpu(), qz.cpu(), qs.cpu(), 128, 2)
    _check_gemv(out, ref, x, dq)


def test_gemv_w2_group64_generic_and_errors():
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import awq_inference_engine as E
    (qw, qz, qs), dq = _packed_layer(128, 1024, 44, G=64)
    x = torch.randn(2, 1024, generator=torch.Generator().manual_seed(1)).to(torch.float16).to(dev)
    out = E.gemv_forward_cuda(x, qw, qs, qz, 2, 64)
    ref = oracle.wq_matmul(x.cpu(), qw.cpu(), qz.cpu(), qs.cpu(), 64, 2)
    _check_gemv(out, ref, x, dq)
    with pytest.raises(NotImplementedError):
        E.gemv_forward_cuda(x, qw, qs, qz, 4, 64)
    with pytest.raises(TypeError):
        E.gemv_forward_cuda(x.float(), qw, qs, qz, 2, 64)
    with pytest.raises(L.BdqError):
        E.gemv_forward_cuda(x.cpu(), qw, qs, qz, 2, 64)


@pytest.mark.parametrize("oc,ic", [(12288, 4096), (11008, 4096), (48, 2176), (16, 128)])
@pytest.mark.parametrize("flags", [0, 1, 16])
def test_gemv_silu_mul_fused(oc, ic, flags):
    """Fused gate/up launch with the SiLU-and-multiply epilogue (fused_mlp.py:44-108): equal, up to the last fp16 ulp of
    the sigmoid, to the two stand-alone GEMVs followed by F.silu and the product (same kernel arithmetic per matrix)."""
    from bitdistiller_b200 import awq_inference_engine as E
    (gw, gz, gs), _ = _packed_layer(oc, ic, 300)
    (uw, uz, us), _ = _packed_layer(oc, ic, 301)
    x = torch.randn(1, ic, generator=torch.Generator().manual_seed(11)).to(torch.float16).to(dev)
    out = E.gemv_silu_mul_forward_cuda(x, (gw, gs, gz), (uw, us, uz), 128, flags=flags)
    g1 = E.gemv_forward_cuda(x, gw, gs, gz, 2, 128)
    u1 = E.gemv_forward_cuda(x, uw, us, uz, 2, 128)
    ref = torch.nn.functional.silu(g1) * u1
    assert out.shape == (1, oc) and out.dtype == torch.float16
    err = (out.float() - ref.float()).abs()
    tol = 2.0 ** -9 * ref.float().abs() + 2.0 ** -24   # 2 fp16 ulps
    assert not bool((err > tol).any()), (int((err > tol).sum()), float(err.max()))
    assert float((out != ref).float().mean()) < 0.02   # and almost always bit-equal
    if oc > 16:
        with pytest.raises(ValueError):
            E.gemv_silu_mul_forward_cuda(x, (gw, gs, gz), (uw[:16], us[:16], uz[:16]), 128)


def test_quant_llama_mlp_matches_dequant():
    """QuantLlamaMLP / make_fused_mlp (inference/modules/fused_mlp.py:18-133) for 1, 4 and 64 rows against the
    dequantised fp32 MLP."""
    from bitdistiller_b200.fused_mlp import QuantLlamaMLP, make_fused_mlp
    from bitdistiller_b200.qmodule import WQLinear
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    H, I = 1024, 2816

    class MLP(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.gate_proj = torch.nn.Linear(H, I, bias=False, dtype=torch.float16, device=dev)
            self.up_proj = torch.nn.Linear(H, I, bias=False, dtype=torch.float16, device=dev)
            self.down_proj = torch.nn.Linear(I, H, bias=False, dtype=torch.float16, device=dev)

    class Block(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.mlp = MLP()

    blk = Block()
    dq = {}
    for i, name in enumerate(["gate_proj", "up_proj", "down_proj"]):
        lin = getattr(blk.mlp, name)
        lin.weight.data = make_weight("randn", lin.out_features, lin.in_features, torch.float16, 40 + i).to(dev) * (
            (lin.in_features ** -0.5) / 0.02)
        lin.weight.data, s, z = pseudo_quantize_tensor(lin.weight.data, n_bit=2, q_group_size=128, get_scale_zp=True)
        dq[name] = lin.weight.data.float()
        setattr(blk.mlp, name, WQLinear.from_linear(lin, 2, 128, False, s, z))
    make_fused_mlp(blk)
    assert isinstance(blk.mlp, QuantLlamaMLP)
    assert {"gate_proj_qweight", "gate_proj_scales", "gate_proj_qzeros", "up_proj_qweight", "up_proj_scales",
            "up_proj_qzeros"} <= set(dict(blk.mlp.named_buffers()))
    for rows in (1, 4, 64):
        x = torch.randn(1, rows, H, device=dev, dtype=torch.float16)
        y = blk.mlp(x)
        xf = x.float()
        ref = (torch.nn.functional.silu(xf @ dq["gate_proj"].T) * (xf @ dq["up_proj"].T)) @ dq["down_proj"].T
        assert y.shape == (1, rows, H)
        assert float((y.float() - ref).norm() / ref.norm()) < 5e-3, rows


def test_wqlinear_forward_decode_matches_dequant():
    """WQLinear end to end: real_quantize_model_weight-style construction, forward with 1 row and with bias."""
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    from bitdistiller_b200.qmodule import WQLinear
    lin = torch.nn.Linear(4096, 1024, bias=True, dtype=torch.float16, device=dev)
    lin.weight.data = make_weight("randn", 1024, 4096, torch.float16, 3).to(dev)
    lin.weight.data, s, z = pseudo_quantize_tensor(lin.weight.data, n_bit=2, q_group_size=128, get_scale_zp=True)
    ql = WQLinear.from_linear(lin, 2, 128, False, s, z)
    x = torch.randn(1, 1, 4096, device=dev, dtype=torch.float16)
    y = ql(x)
    ref = torch.nn.functional.linear(x.float(), lin.weight.data.float(), lin.bias.float())
    assert y.shape == (1, 1, 1024)
    assert float((y.float() - ref).norm() / ref.norm()) < 3e-3
    sd = ql.state_dict()
    assert set(sd) == {"qweight", "qzeros", "scales", "bias"}
    with pytest.raises(NotImplementedError):
        WQLinear(3, 128, 128, 128, False, dev)


def test_tp_epilogues_write_own_and_peer_slices():
    """bdq_gemv_w2_tp / bdq_gemm_w2_tp on one GPU with a second local buffer standing in for the peer: both buffers
    receive the rank's channel slice at the right column offset and equal the plain kernels' outputs bit for bit."""
    import ctypes
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import awq_inference_engine as E
    from bitdistiller_b200.triton_kernels import gemm_w2
    lib = L.lib()
    (qw, qz, qs), _ = _packed_layer(1024, 4096, 55)
    half = 512                                   # "rank 1 of 2": channels [512, 1024)
    sq, sz, ss = qw[half:].contiguous(), qz[half:].contiguous(), qs[half:].contiguous()
    st = L.stream_ptr(dev)
    # decode
    x = torch.randn(1, 4096, generator=torch.Generator().manual_seed(2)).half().to(dev)
    ref = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128)
    mine = torch.zeros(1, 1024, dtype=torch.float16, device=dev)
    peer = torch.zeros(1, 1024, dtype=torch.float16, device=dev)
    arr = lambda *ps: (ctypes.c_void_p * len(ps))(*ps)
    rc = lib.bdq_gemv_w2_tp(L.ptr(x), 1, arr(sq.data_ptr()), arr(ss.data_ptr()), arr(sz.data_ptr()),
                            arr(mine.data_ptr() + half * 2), (ctypes.c_int64 * 1)(512), L.i64(4096), 128, 0, 1,
                            arr(peer.data_ptr() + half * 2), st)
    L.check(rc, "bdq_gemv_w2_tp")
    assert torch.equal(mine[:, half:], ref[:, half:]) and torch.equal(peer[:, half:], ref[:, half:])
    assert not mine[:, :half].any() and not peer[:, :half].any()
    # prefill
    xm = torch.randn(192, 4096, generator=torch.Generator().manual_seed(3)).half().to(dev)
    refm = gemm_w2(xm, qw, qs, qz, 128)
    mine = torch.zeros(192, 1024, dtype=torch.float16, device=dev)
    peer = torch.zeros(192, 1024, dtype=torch.float16, device=dev)
    rc = lib.bdq_gemm_w2_tp(L.ptr(xm), L.ptr(sq), L.ptr(ss), L.ptr(sz), ctypes.c_void_p(mine.data_ptr() + half * 2),
                            L.i64(192), L.i64(4096), L.i64(512), 128, L.i64(1024), 1, arr(peer.data_ptr() + half * 2), st)
    L.check(rc, "bdq_gemm_w2_tp")
    assert torch.equal(mine[:, half:], refm[:, half:]) and torch.equal(peer[:, half:], refm[:, half:])
    assert not mine[:, :half].any() and not peer[:, :half].any()


def _chain_layers(hidden, kv, inter, layers, silu, seed=0):
    """A dependent decode chain: q/k/v <- x, o <- q, gate/up <- o, down <- gate (or silu(gate)*up), next layer <- down."""
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq

    def packed(oc, ic, s):
        g = torch.Generator(device=dev).manual_seed(s)
        w = (torch.randn(oc, ic, device=dev, generator=g) * ic ** -0.5).half()
        dq, sc, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
        qw, qz, qs = pack_wq(dq, sc.view(oc, -1), z.view(oc, -1), 2, 128)
        return qw, qs, qz, torch.zeros(1, oc, dtype=torch.float16, device=dev)

    x0 = torch.randn(1, hidden, generator=torch.Generator().manual_seed(seed)).half().to(dev)
    phases, x = [], x0
    for li in range(layers):
        q, k, v = packed(hidden, hidden, 10 * li + 1), packed(kv, hidden, 10 * li + 2), packed(kv, hidden, 10 * li + 3)
        o = packed(hidden, hidden, 10 * li + 4)
        gt, up = packed(inter, hidden, 10 * li + 5), packed(inter, hidden, 10 * li + 6)
        dn = packed(hidden, inter, 10 * li + 7)
        phases.append({"x": x, "mats": [q, k, v], "nowait": li == 0})
        phases.append({"x": q[3], "mats": [o]})
        if silu:
            phases.append({"x": o[3], "mats": [gt, (up[0], up[1], up[2], None)], "silu": True})
        else:
            phases.append({"x": o[3], "mats": [gt, up]})
        phases.append({"x": gt[3], "mats": [dn]})
        x = dn[3]
    return phases


@pytest.mark.parametrize("hidden,kv,inter,layers,silu", [(1024, 256, 2816, 3, False), (4096, 1024, 12288, 2, False),
                                                         (512, 128, 1408, 4, True), (4096, 1024, 11008, 1, True)])
def test_decode_chain_equals_launch_sequence(hidden, kv, inter, layers, silu):
    """bdq_gemv_w2_chain_* (one persistent launch, grid barriers between dependent phases) is bit-equal to the same phases
    issued as consecutive bdq_gemv_w2_multi launches; a second run of the same plan (counters reset in-kernel) too."""
    from bitdistiller_b200 import awq_inference_engine as E
    phases = _chain_layers(hidden, kv, inter, layers, silu)

    def outs():
        return [m[3] for ph in phases for m in ph["mats"] if m[3] is not None]

    for ph in phases:   # the reference sequence: one launch per phase
        mats = ph["mats"]
        if ph.get("silu"):
            E.gemv_silu_mul_forward_cuda(ph["x"], mats[0][:3], mats[1][:3], 128, out=mats[0][3])
        else:
            E.gemv_multi_forward_cuda(ph["x"], [m[0] for m in mats], [m[1] for m in mats], [m[2] for m in mats], 128,
                                      outs=[m[3] for m in mats])
    torch.cuda.synchronize()
    ref = [o.clone() for o in outs()]
    assert all(bool(r.isfinite().all()) and float(r.abs().max()) > 0 for r in ref)
    for cap in (None, 5):    # 5: the chain split over several launches (a launch holds a bounded number of phases)
        chain = E.DecodeChain(phases, max_phases_per_launch=cap)
        for rep in range(3):
            for o in outs():
                o.zero_()
            chain.run()
            torch.cuda.synchronize()
            for i, (a, b) in enumerate(zip(outs(), ref)):
                assert torch.equal(a, b), (cap, rep, i, float((a.float() - b.float()).abs().max()))


def test_decode_chain_errors():
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import awq_inference_engine as E
    phases = _chain_layers(512, 128, 1408, 1, False)
    bad = [dict(phases[0], x=torch.randn(2, 512, device=dev).half())]
    with pytest.raises(ValueError):
        E.DecodeChain(bad)
    with pytest.raises(L.BdqError):
        E.DecodeChain([dict(phases[0], x=torch.randn(1, 512).half())])
