"""GPU parity tests for the fused fake-quant / STE kernels (through the C ABI via bitdistiller_b200.quantizer).

Bars (BASELINE.json north_star): codes, zeros, scales bit-exact; dequantised weights bit-exact here (the stated
bar is <= 1 bf16 ulp); backward within the reference's own T-dtype autograd noise of the float64 closed form.
"""
import os

import numpy as np
import pytest
import torch

import oracle
from _util import (DT, GOLDEN, bits_equal, load_cases, make_weight, t, torch_int_fakequant, torch_nf3_fakequant,
                   ulp_of)

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _cpu_semantics(monkeypatch):
    monkeypatch.setenv("BDQ_SEMANTICS", "cpu")


def dev():
    return torch.device("cuda:0")


def test_pseudo_quantize_tensor_golden_bit_exact():
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    n = 0
    for meta, c in load_cases("pq_int"):
        dname, kind, n_bit, G = meta[0], meta[1], int(meta[2]), int(meta[3])
        dt = DT[dname]
        w = t(c["w"], dt).to(dev())
        dq, s, z = pseudo_quantize_tensor(w, n_bit=n_bit, zero_point=True, q_group_size=G, get_scale_zp=True)
        assert dq.dtype == dt and dq.shape == w.shape
        assert bits_equal(dq.cpu(), t(c["dq"], dt)), meta
        assert bits_equal(s.cpu(), t(c["s"], dt)), meta
        assert bits_equal(z.cpu(), t(c["z"], dt)), meta
        w2 = w.clone()
        r = pseudo_quantize_tensor(w2, n_bit=n_bit, zero_point=True, q_group_size=G, inplace=True)
        assert bits_equal(w2.cpu(), t(c["dq"], dt)) and r.data_ptr() == w2.data_ptr(), meta
        n += 1
    assert n == 105


def _bwd_check(gw, w, gy, G, ref64, dt, meta):
    """|ours - closed form(double)| bounded by the rounding of the output dtype plus fp32 accumulation noise."""
    gw = gw.double().cpu()
    # elements equal to their group's max / min also receive the amax / amin (scale-path) gradient: the packed kernels
    # add it as a hi + lo pair of T values (two roundings, <= 1 ulp each way); the reference's own T-dtype autograd
    # accumulates these terms with three roundings.
    wg = w.double().reshape(-1, G)
    ext = ((wg == wg.amax(1, keepdim=True)) | (wg == wg.amin(1, keepdim=True))).reshape(ref64.shape)
    tol = torch.where(ext, 2.01, 1.01) * ulp_of(ref64, dt) + 1e-30
    gsum = gy.double().abs().reshape(-1, G).sum(1, keepdim=True).expand(-1, G).reshape(ref64.shape)
    tol = tol + 1e-5 * gsum
    bad = (gw - ref64).abs() > tol
    assert not bool(bad.any()), (meta, int(bad.sum()), float((gw - ref64).abs().max()))


def test_ste_quantizers_golden():
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, SteInt3AsymQuantizer, SteN2F3Quantizer
    n = 0
    for meta, c in load_cases("ste"):
        qname, dname, kind, G = meta[0], meta[1], meta[2], int(meta[3])
        dt = DT[dname]
        w, gy = t(c["w"], dt), t(c["gy"], dt)
        mod = {"int2": SteInt2AsymQuantizer, "int3": SteInt3AsymQuantizer, "nf3": SteN2F3Quantizer}[qname](G)
        x = w.to(dev()).requires_grad_(True)
        y = mod(x)
        y.backward(gy.to(dev()))
        assert bits_equal(y.detach().cpu(), t(c["y"], dt)), meta
        if qname == "nf3":
            ref64 = oracle.nf3_ste_bwd(w, gy, G)
        else:
            ref64 = oracle.int_ste_bwd(w, gy, G, int(qname[-1]))
        _bwd_check(x.grad, w, gy, G, ref64, dt, meta)
        n += 1
    assert n == 90
    z = np.load(os.path.join(GOLDEN, "ste.npz"))
    from bitdistiller_b200.quantizer import pseudo_quantize_n2f3_tensor
    y = pseudo_quantize_n2f3_tensor(t(z["nan_w"], torch.bfloat16).to(dev()), 128)
    assert bits_equal(y.cpu(), t(z["nan_y"], torch.bfloat16))


def _ref_autograd_check(gw, c, w, gy, G, dt, meta):
    """The CUDA backward against the REFERENCE's own autograd results stored in ste.npz (make_golden.py: `gw` = the
    reference run in dtype T, `gw64` / `y64` = the same module run in float64), not only against the oracle.
    (1) vs `gw`: within the T-run's own rounding noise (it accumulates the scale-path sum in T) -- the bound the oracle
        itself is held to in test_oracle_golden.py.
    (2) vs `gw64`: in every group where the float64 run took the same codes as the T run (|y - y64| far below one
        quantisation step), within 2 ulp of T plus the effect of the T-rounded scale on (q - w/s): 1e-7 (fp32) /
        1e-2 (16-bit) of sum|gy| of the group.  For fp32 this is a tight pin on the closed form."""
    gw = gw.double().cpu().reshape(-1, G)
    gs = gy.double().abs().reshape(-1, G).sum(1, keepdim=True)
    ref_T = t(c["gw"], torch.float64).reshape(-1, G)
    tol = 8 * ulp_of(ref_T, dt) + 1e-6 if dt != torch.float32 else 1e-4 * (ref_T.abs() + 1.0)
    noise = {torch.float32: 1e-5, torch.bfloat16: 4e-2, torch.float16: 5e-3}[dt]
    bad = (gw - ref_T).abs() > tol + noise * gs
    assert bad.float().mean() < 0.002, (meta, "vs reference T autograd", float(bad.float().mean()))
    y, y64 = t(c["y"], torch.float64).reshape(-1, G), torch.from_numpy(c["y64"]).reshape(-1, G)
    wg = w.double().reshape(-1, G)
    agree = ((y - y64).abs() <= 0.04 * (wg.amax(1, keepdim=True) - wg.amin(1, keepdim=True)) + 1e-12).all(1)
    g64 = torch.from_numpy(c["gw64"]).reshape(-1, G)
    noise64 = {torch.float32: 1e-7, torch.bfloat16: 1e-2, torch.float16: 1e-2}[dt]
    bad = ((gw - g64).abs() > 2.01 * ulp_of(g64, dt) + noise64 * gs + 1e-30)[agree]
    assert not bool(bad.any()), (meta, "vs reference f64 autograd", int(bad.sum()))
    return int(agree.sum()), int(agree.numel())


def test_ste_backward_vs_reference_autograd():
    """the stored reference gradients are read by a GPU test."""
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, SteInt3AsymQuantizer, SteN2F3Quantizer
    n = ag = tot = 0
    for meta, c in load_cases("ste"):
        qname, dname, kind, G = meta[0], meta[1], meta[2], int(meta[3])
        dt = DT[dname]
        w, gy = t(c["w"], dt), t(c["gy"], dt)
        mod = {"int2": SteInt2AsymQuantizer, "int3": SteInt3AsymQuantizer, "nf3": SteN2F3Quantizer}[qname](G)
        x = w.to(dev()).requires_grad_(True)
        mod(x).backward(gy.to(dev()))
        a, b = _ref_autograd_check(x.grad, c, w, gy, G, dt, meta)
        ag, tot, n = ag + a, tot + b, n + 1
    assert n == 90 and ag > 0.7 * tot  # most groups are checked against the float64 autograd too


@pytest.mark.parametrize("dname", ["f32", "bf16", "f16"])
def test_backward_cuda_semantics(dname, monkeypatch):
    """the backward under BDQ_SEMANTICS=cuda (the package default), against the oracle's sem=1 closed
    form AND against eager-PyTorch autograd of the reference formula run on the GPU itself in float64 where codes agree."""
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, SteInt3AsymQuantizer, SteN2F3Quantizer
    monkeypatch.setenv("BDQ_SEMANTICS", "cuda")
    dt = DT[dname]
    w = make_weight("randn", 128, 2048, dt, 17)
    lim = w.float().reshape(128, -1, 128).abs().amax(-1, keepdim=True) * 0.8
    w = torch.clamp(w.float().reshape(128, -1, 128), -lim, lim).reshape(128, 2048).to(dt)
    gy = make_weight("uniform", 128, 2048, dt, 18)
    for mod, orc in [(SteInt2AsymQuantizer(128), lambda: oracle.int_ste_bwd(w, gy, 128, 2, sem=1)),
                     (SteInt3AsymQuantizer(128), lambda: oracle.int_ste_bwd(w, gy, 128, 3, sem=1)),
                     (SteInt2AsymQuantizer(64), lambda: oracle.int_ste_bwd(w, gy, 64, 2, sem=1)),
                     (SteN2F3Quantizer(128), lambda: oracle.nf3_ste_bwd(w, gy, 128, sem=1))]:
        x = w.to(dev()).requires_grad_(True)
        y = mod(x)
        y.backward(gy.to(dev()))
        _bwd_check(x.grad, w, gy, mod.q_group_size, orc(), dt, (dname, type(mod).__name__, "sem=cuda"))
        if isinstance(mod, SteN2F3Quantizer):
            assert bits_equal(y.detach().cpu(), torch_nf3_fakequant(w.to(dev()), 128).cpu())
        else:
            e_dq, _, _ = torch_int_fakequant(w.to(dev()), mod.bit, mod.q_group_size, True)
            assert bits_equal(y.detach().cpu(), e_dq.cpu()), (dname, type(mod).__name__)


def test_bench_launch_shape_vs_oracle():
    """the launch shape bench.py times -- ONE launch over the flat 202.4 M-weight bf16 buffer of a
    Llama-2-7B decoder layer (whole-wave grid) -- forward and STE backward against the oracle on sampled slabs
    (first / last / wave boundaries / random), bit-exact forward, backward within the closed-form bound."""
    from bitdistiller_b200 import _lib as L
    G, n = 128, 4 * 4096 * 4096 + 3 * 11008 * 4096
    gen = torch.Generator(device=dev()).manual_seed(5)
    w = (torch.randn(n, device=dev(), generator=gen) * 0.02).to(torch.bfloat16)
    gy = torch.randn(n, device=dev(), generator=gen).to(torch.bfloat16)
    dq, gw = torch.empty_like(w), torch.empty_like(w)
    lib, st = L.lib(), L.stream_ptr(dev())
    L.check(lib.bdq_int_fakequant_fwd(L.ptr(w), L.ptr(dq), None, None, None, None, None, L.i64(n), G, 2, L.BF16,
                                      L.ROUND_HALF_AWAY, L.SEM_CPU, st), "fwd")
    L.check(lib.bdq_int_ste_bwd(L.ptr(w), L.ptr(gy), L.ptr(gw), L.i64(n), G, 2, L.BF16, L.SEM_CPU, st), "bwd")
    slab = 64 * 1024
    rs = np.random.RandomState(0)
    starts = [0, n - slab, n // 2 - slab // 2] + [int(s) * G for s in rs.randint(0, (n - slab) // G, size=13)]
    for s0 in starts:
        ws, gs = w[s0:s0 + slab].cpu(), gy[s0:s0 + slab].cpu()
        ref = oracle.int_fakequant(ws, G, 2, round_mode=1)
        assert bits_equal(dq[s0:s0 + slab].cpu(), ref["dq"].reshape(-1)), s0
        _bwd_check(gw[s0:s0 + slab], ws, gs, G, oracle.int_ste_bwd(ws, gs, G, 2).reshape(-1), torch.bfloat16, s0)
    # checksum of the whole buffer against a second, differently shaped launch sequence (per 16 M-element chunk)
    dq2 = torch.empty_like(w)
    step = 16 * 1024 * 1024
    for s0 in range(0, n, step):
        m = min(step, n - s0)
        L.check(lib.bdq_int_fakequant_fwd(L.ptr(w[s0:]), L.ptr(dq2[s0:]), None, None, None, None, None, L.i64(m), G, 2,
                                          L.BF16, L.ROUND_HALF_AWAY, L.SEM_CPU, st), "fwd chunk")
    assert torch.equal(dq.view(torch.int16), dq2.view(torch.int16))


@pytest.mark.parametrize("dname", ["f32", "bf16", "f16"])
@pytest.mark.parametrize("kind", ["randn", "uniform", "heavy"])
def test_large_vs_oracle(dname, kind):
    """[512, 4096] per dtype: codes / scales / zeros / dq bit-exact against the oracle, both rounding modes, + NF3."""
    from bitdistiller_b200 import quantizer as Q
    dt = DT[dname]
    w = make_weight(kind, 512, 4096, dt, 1234)
    wd = w.to(dev())
    for bits, G, rm in [(2, 128, 0), (2, 128, 1), (2, 64, 1), (3, 128, 1), (4, 32, 0), (2, 4096, 0)]:
        ref = oracle.int_fakequant(w, G, bits, round_mode=rm)
        dq, s, z, codes = Q.int_fakequant(wd, bits, G, rm, want_scale_zp=True, want_codes=True)
        assert bits_equal(dq.cpu(), ref["dq"]), (bits, G, rm)
        assert bits_equal(s.cpu().reshape(ref["scales"].shape), ref["scales"])
        assert bits_equal(z.cpu().reshape(ref["zeros"].shape), ref["zeros"])
        assert torch.equal(codes.cpu(), ref["codes"])
    for G in [128, 64]:
        ref = oracle.nf3_fakequant(w, G)
        dq, sp, sn, codes = Q.nf3_fakequant(wd, G, want_codes=True, want_scales=True)
        assert bits_equal(dq.cpu(), ref["dq"])
        assert torch.equal(codes.cpu(), ref["codes"])
        assert bits_equal(sp.cpu(), ref["scale_pos"]) and bits_equal(sn.cpu(), ref["scale_neg"])


@pytest.mark.parametrize("dname", ["f32", "bf16"])
def test_large_backward_vs_oracle(dname):
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, SteN2F3Quantizer
    dt = DT[dname]
    w = make_weight("randn", 256, 4096, dt, 7)
    lim = w.float().reshape(256, -1, 128).abs().amax(-1, keepdim=True) * 0.8
    w = torch.clamp(w.float().reshape(256, -1, 128), -lim, lim).reshape(256, 4096).to(dt)  # ties, as after apply_clip
    gy = make_weight("uniform", 256, 4096, dt, 8)
    for mod, orc in [(SteInt2AsymQuantizer(128), lambda: oracle.int_ste_bwd(w, gy, 128, 2)),
                     (SteInt2AsymQuantizer(-1), lambda: oracle.int_ste_bwd(w, gy, 4096, 2)),
                     (SteN2F3Quantizer(128), lambda: oracle.nf3_ste_bwd(w, gy, 128))]:
        x = w.to(dev()).requires_grad_(True)
        mod(x).backward(gy.to(dev()))
        G = mod.q_group_size if mod.q_group_size > 0 else 4096
        _bwd_check(x.grad, w, gy, G, orc(), dt, (dname, type(mod).__name__))


def test_fused_clip_matches_apply_clip_then_quant():
    from bitdistiller_b200 import quantizer as Q
    for dname in ["bf16", "f32"]:
        dt = DT[dname]
        w = make_weight("randn", 128, 1024, dt, 3)
        g = torch.Generator().manual_seed(5)
        mx = w.float().reshape(128, -1, 128).amax(-1)
        mn = w.float().reshape(128, -1, 128).amin(-1)
        cmax = (mx * (1 - torch.randint(0, 10, mx.shape, generator=g) / 20)).to(dt)
        cmin = (mn * (1 - torch.randint(0, 10, mn.shape, generator=g) / 20)).to(dt)
        ref_i = oracle.int_fakequant(w, 128, 2, round_mode=1, clip_min=cmin, clip_max=cmax)
        ref_n = oracle.nf3_fakequant(w, 128, clip_min=cmin, clip_max=cmax)
        wd, a, b = w.to(dev()), cmin.to(dev()), cmax.to(dev())
        dq, _, _, codes = Q.int_fakequant(wd, 2, 128, 1, want_codes=True, clip_min=a, clip_max=b)
        assert bits_equal(dq.cpu(), ref_i["dq"]) and torch.equal(codes.cpu(), ref_i["codes"])
        dqn, _, _, _ = Q.nf3_fakequant(wd, 128, clip_min=a, clip_max=b)
        assert bits_equal(dqn.cpu(), ref_n["dq"])


def test_cuda_eager_semantics_hypothesis(monkeypatch):
    """BDQ_SEMANTICS=cuda must reproduce eager PyTorch ON THE GPU bit for bit, and the
    CPU / CUDA eager results must differ exactly where the oracle's sem=0 / sem=1 differ."""
    from bitdistiller_b200 import quantizer as Q
    monkeypatch.setenv("BDQ_SEMANTICS", "cuda")
    for dname in ["f32", "bf16", "f16"]:
        dt = DT[dname]
        w = make_weight("randn", 256, 2048, dt, 99)
        wd = w.to(dev())
        for bits, G, rm in [(2, 128, 0), (2, 64, 1), (4, 128, 0)]:
            e_dq, e_s, e_z = torch_int_fakequant(wd, bits, G, bool(rm))
            dq, s, z, _ = Q.int_fakequant(wd, bits, G, rm, want_scale_zp=True)
            assert bits_equal(s.cpu().reshape(e_s.shape), e_s.cpu()), (dname, bits, G, rm, "scales")
            assert bits_equal(z.cpu().reshape(e_z.shape), e_z.cpu()), (dname, bits, G, rm, "zeros")
            assert bits_equal(dq.cpu(), e_dq.cpu()), (dname, bits, G, rm, "dq")
            ref1 = oracle.int_fakequant(w, G, bits, round_mode=rm, sem=1)
            assert bits_equal(dq.cpu(), ref1["dq"])
        e = torch_nf3_fakequant(wd, 128)
        dq, _, _, _ = Q.nf3_fakequant(wd, 128)
        assert bits_equal(dq.cpu(), e.cpu()), (dname, "nf3")


def test_errors_and_edge_cases():
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, pseudo_quantize_tensor
    with pytest.raises(AssertionError):
        pseudo_quantize_tensor(torch.zeros(4, 100, device=dev()), n_bit=2, q_group_size=128)
    with pytest.raises(L.BdqError):
        pseudo_quantize_tensor(torch.zeros(4, 128), n_bit=2, q_group_size=128)  # CPU tensor: no fallback
    with pytest.raises(TypeError):
        pseudo_quantize_tensor(torch.zeros(4, 128, device=dev(), dtype=torch.float64), n_bit=2, q_group_size=128)
    e = pseudo_quantize_tensor(torch.zeros(0, 128, device=dev()), n_bit=2, q_group_size=128)
    assert e.shape == (0, 128)
    # 4-D input as autoclip passes it (autoclip.py:28,56), ragged tail (rows not a multiple of the warp chunk)
    w = make_weight("randn", 3, 5 * 128, torch.bfloat16, 1).reshape(3, 1, 5, 128)
    dq = pseudo_quantize_tensor(w.to(dev()), n_bit=2, q_group_size=128)
    assert dq.shape == w.shape and bits_equal(dq.cpu(), oracle.int_fakequant(w, 128, 2)["dq"])
    # non-contiguous view and unaligned base pointer
    base = make_weight("randn", 8, 260, torch.float32, 2).to(dev())
    v = base[:, 4:]
    dq = pseudo_quantize_tensor(v, n_bit=2, q_group_size=128)
    assert bits_equal(dq.cpu(), oracle.int_fakequant(v.cpu().contiguous(), 128, 2)["dq"])
    flat = make_weight("randn", 1, 128 * 9 + 1, torch.float32, 4).to(dev()).reshape(-1)[1:].reshape(9, 128)
    dq = pseudo_quantize_tensor(flat, n_bit=2, q_group_size=128)
    assert bits_equal(dq.cpu(), oracle.int_fakequant(flat.cpu(), 128, 2)["dq"])
    q = SteInt2AsymQuantizer()
    assert q.q_group_size == 64 and q.bit == 2


@pytest.mark.parametrize("G", [128, 64, 32])
@pytest.mark.parametrize("scale", [1.0, 1e-12, 1e-33, 2e-37, 1e31, 1e36])
def test_nf3_fp32_lean_forward_vs_oracle(scale, G):
    """The fp32 NF3 forward without auxiliary outputs (shared-divisor IEEE division, one branch-free staircase) against the CPU
    oracle, bit for bit -- including groups whose scales leave the fast path's exponent range (per-element __fdiv_rn), denormal
    weights, exact zeros, a group with max == 0 (NaN by the reference's 0/0) and the fused clip."""
    from bitdistiller_b200 import quantizer as Q
    w = make_weight("heavy", 96, 1024, torch.float32, 21) * (scale / 0.02)
    w[3, :G] = w[3, :G].abs() * -1.0          # an all-negative group: max < 0, both scales from the same sign
    w[5, 7] = 0.0
    w[6, :G] = torch.minimum(w[6, :G], torch.zeros(G))
    w[6, 0] = 0.0                              # max == 0 exactly -> NaN group
    w[7, G:2 * G] = 0.0                        # all-zero group -> NaN group
    dq = Q.nf3_fakequant(w.to(dev()), G)[0]
    ref = oracle.nf3_fakequant(w, G)
    assert bits_equal(dq.cpu(), ref["dq"]), (scale, G)
    cmax = (w.view(-1, G).amax(1) * 0.7).contiguous()
    cmin = (w.view(-1, G).amin(1) * 0.8).contiguous()
    dq = Q.nf3_fakequant(w.to(dev()), G, clip_min=cmin.to(dev()), clip_max=cmax.to(dev()))[0]
    ref = oracle.nf3_fakequant(w, G, clip_min=cmin, clip_max=cmax)
    assert bits_equal(dq.cpu(), ref["dq"]), (scale, G, "clip")


@pytest.mark.parametrize("case", range(6))
def test_ste_adversarial_groups(case):
    """tools/stress_ste_bwd.py on a few seeds: groups far from zero (clamp masks of the threshold backward), constant and
    two-valued groups, ties at the extremes, tiny / huge ranges -- forward bit for bit, backward within _bwd_check's bound,
    for int2 / int3 / nf3 and groups 128 / 64 / 32."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import stress_ste_bwd as S
    from bitdistiller_b200.quantizer import SteInt2AsymQuantizer, SteInt3AsymQuantizer, SteN2F3Quantizer
    g = torch.Generator().manual_seed(900 + case)
    dt = [torch.bfloat16, torch.float16][case % 2]
    G = [128, 64, 32][case % 3]
    w = S.adversarial(96, 1024, G, dt, g)
    gy = torch.randn(96, 1024, generator=g).to(dt)
    for name, mod, ob, fr in (
            ("int2", SteInt2AsymQuantizer(G), lambda a, b: oracle.int_ste_bwd(a, b, G, 2), lambda a: oracle.int_fakequant(a, G, 2, 1)["dq"]),
            ("int3", SteInt3AsymQuantizer(G), lambda a, b: oracle.int_ste_bwd(a, b, G, 3), lambda a: oracle.int_fakequant(a, G, 3, 1)["dq"]),
            ("nf3", SteN2F3Quantizer(G), lambda a, b: oracle.nf3_ste_bwd(a, b, G), lambda a: oracle.nf3_fakequant(a, G)["dq"])):
        assert S.check(name, mod, ob, w, gy, G, dt, fr) == 0, (name, dt, G)
