"""Parity pins against the UNMODIFIED reference running ON THE B200: the reference's Python in CUDA
eager (quantizer.py, qlinear.py, qmodule.py), its Triton kernel `quant_matmul_v2` (triton_kernels.py:14-134) and its CUDA
`gemv_kernel_g128_2bit` (gemv_cuda.cu:202-272) compiled for sm_100 -- all shipped under baseline/_ref/ by
baseline/build_ref.py.  The package runs with its default BDQ_SEMANTICS=cuda here (the reference's own device).

Bars (BASELINE.json north_star): codes / zeros / packed words bit-exact; dequantised weights bit-exact (stated bar: 1 bf16
ulp); GEMV / GEMM outputs within rtol 1e-2 of the reference kernels."""
import numpy as np
import pytest
import torch

import oracle
from _ref import load
from _util import DT, QL_GRADS, GOLDEN, bits_equal, make_weight, tiny_llama, ulp_of

pytestmark = pytest.mark.gpu
dev = torch.device("cuda:0")


@pytest.fixture(autouse=True)
def _default_semantics(cuda_semantics):
    """conftest's `cuda_semantics`: BDQ_SEMANTICS=cuda (the package default)."""


@pytest.fixture(scope="module")
def ref():
    return load()


def _weights(dt, rows=64, cols=1024):
    out = [make_weight(k, rows, cols, dt, 40 + i) for i, k in enumerate(["randn", "uniform", "heavy"])]
    w = make_weight("randn", rows, cols, dt, 50)
    lim = w.float().reshape(rows, -1, 128).abs().amax(-1, keepdim=True) * 0.75   # ties at the extremes, as after apply_clip
    out.append(torch.clamp(w.float().reshape(rows, -1, 128), -lim, lim).reshape(rows, cols).to(dt))
    e = make_weight("randn", rows, cols, dt, 51)
    e[0] = 0.0
    e[1] = 0.0123
    e[2] = e[2] * 1e-5
    e[3] = e[3].abs()
    e[4] = -e[4].abs()
    out.append(e)
    return out


# ---------------------------------------------------------------- a1-a4: the quantisers in CUDA eager
@pytest.mark.parametrize("dname", ["f32", "bf16", "f16"])
def test_pseudo_quantize_tensor_vs_reference_cuda_eager(ref, dname):
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    dt, n = DT[dname], 0
    for w in _weights(dt):
        wd = w.to(dev)
        for n_bit, G in [(2, 128), (2, 64), (3, 128), (4, 128), (8, 128), (2, -1), (4, 32)]:
            r_dq, r_s, r_z = ref["quantizer"].pseudo_quantize_tensor(wd.clone(), n_bit=n_bit, zero_point=True,
                                                                     q_group_size=G, get_scale_zp=True)
            dq, s, z = pseudo_quantize_tensor(wd.clone(), n_bit=n_bit, zero_point=True, q_group_size=G, get_scale_zp=True)
            assert bits_equal(s.cpu(), r_s.cpu()), (dname, n_bit, G, "scales")
            assert bits_equal(z.cpu(), r_z.cpu()), (dname, n_bit, G, "zeros")
            assert bits_equal(dq.cpu(), r_dq.cpu()), (dname, n_bit, G, "dq")
            # and the oracle's sem=1 is the same function
            o = oracle.int_fakequant(w, G if G > 0 else w.shape[1], n_bit, round_mode=0, sem=1)
            assert bits_equal(o["dq"], r_dq.cpu()), (dname, n_bit, G, "oracle sem=1")
            n += 1
    assert n == 35


@pytest.mark.parametrize("dname", ["f32", "bf16", "f16"])
def test_ste_quantizers_vs_reference_cuda_autograd(ref, dname):
    """Forward bit-equal; backward against the reference's own autograd on the GPU: run in dtype T (noise-limited: it
    accumulates the scale-path sums in T) and in float64 (tight wherever the float64 run takes the same codes)."""
    import bitdistiller_b200.quantizer as Q
    dt = DT[dname]
    R = ref["quantizer"]
    n = agree_tot = tot = 0
    for wi, w in enumerate(_weights(dt, 32, 512)):
        gy = make_weight("uniform", 32, 512, dt, 70 + wi)
        for name, G in [("SteInt2AsymQuantizer", 128), ("SteInt2AsymQuantizer", 64), ("SteInt3AsymQuantizer", 128),
                        ("SteN2F3Quantizer", 128), ("SteN2F3Quantizer", 64)]:
            if name == "SteN2F3Quantizer" and wi == 4:
                continue   # all-zero / one-signed groups: NaN by construction in the reference, forward-only
            x = w.to(dev).requires_grad_(True)
            y = getattr(Q, name)(G)(x)
            y.backward(gy.to(dev))
            xr = w.to(dev).requires_grad_(True)
            yr = getattr(R, name)(G)(xr)
            yr.backward(gy.to(dev))
            assert bits_equal(y.detach().cpu(), yr.detach().cpu()), (dname, name, G, wi, "forward")
            x64 = w.to(dev).double().requires_grad_(True)
            y64 = getattr(R, name)(G)(x64)
            y64.backward(gy.to(dev).double())
            gw, gT, g64 = x.grad.double().cpu().reshape(-1, G), xr.grad.double().cpu().reshape(-1, G), x64.grad.cpu().reshape(-1, G)
            gs = gy.double().abs().reshape(-1, G).sum(1, keepdim=True)
            tol = 8 * ulp_of(gT, dt) + 1e-6 if dt != torch.float32 else 1e-4 * (gT.abs() + 1.0)
            noise = {torch.float32: 1e-5, torch.bfloat16: 4e-2, torch.float16: 5e-3}[dt]
            bad = (gw - gT).abs() > tol + noise * gs
            assert bad.float().mean() < 0.002, (dname, name, G, wi, "vs T autograd", float(bad.float().mean()))
            wg = w.double().reshape(-1, G)
            rng = wg.amax(1, keepdim=True) - wg.amin(1, keepdim=True)
            agree = ((yr.detach().double().cpu().reshape(-1, G) - y64.detach().cpu().reshape(-1, G)).abs()
                     <= 0.04 * rng + 1e-12).all(1)
            noise64 = {torch.float32: 1e-7, torch.bfloat16: 4e-2, torch.float16: 1e-2}[dt]
            bad = ((gw - g64).abs() > 2.01 * ulp_of(g64, dt) + noise64 * gs + 1e-30)[agree]
            assert not bool(bad.any()), (dname, name, G, wi, "vs f64 autograd", int(bad.sum()))
            agree_tot, tot, n = agree_tot + int(agree.sum()), tot + agree.numel(), n + 1
    assert n == 23 and agree_tot > 0.6 * tot


def test_nf3_nan_case_vs_reference(ref):
    from bitdistiller_b200.quantizer import pseudo_quantize_n2f3_tensor
    w = make_weight("randn", 4, 256, torch.bfloat16, 77)
    w[0, :128] = w[0, :128].abs()
    w[0, 5] = 0.0
    w[1] = 0.0
    y = pseudo_quantize_n2f3_tensor(w.to(dev), 128)
    yr = ref["quantizer"].pseudo_quantize_n2f3_tensor(w.to(dev), 128)
    assert bits_equal(y.cpu(), yr.cpu()) and bool(yr.isnan().any())


# ---------------------------------------------------------------- a5: QLinear / convertModelToQuant
def _run_model(convert, quant_type, ids):
    model = tiny_llama().to(dev)
    model, replaced = convert(model, compute_dtype=torch.bfloat16, quant_type=quant_type, q_group_size=128)
    assert replaced
    res = model(input_ids=ids, labels=ids)
    res.loss.backward()
    return model, res


@pytest.mark.parametrize("quant_type", ["int2-asym", "ste-n2f3"])
def test_qlinear_convert_vs_reference_on_gpu(ref, quant_type):
    """quantization/qlinear.py:5-78 (what train/train.py:448 calls): the reference's convertModelToQuant + QLinear on a
    tiny Llama in CUDA eager against the drop-in: same module tree, logits bit-equal (the fake-quantised weights are
    bit-equal and both sides call the same cuBLAS), every linear's weight gradient within the reference's own bf16
    autograd noise."""
    from bitdistiller_b200 import qlinear as ours
    ids = torch.randint(0, 128, (2, 32), generator=torch.Generator().manual_seed(11)).to(dev)
    m_ref, r_ref = _run_model(ref["qlinear"].convertModelToQuant, quant_type, ids)
    m_our, r_our = _run_model(ours.convertModelToQuant, quant_type, ids)
    names_ref = [n for n, m in m_ref.named_modules() if isinstance(m, ref["qlinear"].QLinear)]
    names_our = [n for n, m in m_our.named_modules() if isinstance(m, ours.QLinear)]
    assert names_ref == names_our and len(names_our) == 14 and "lm_head" not in names_our
    assert not isinstance(m_our.lm_head, ours.QLinear)
    assert torch.equal(r_our.logits, r_ref.logits), float((r_our.logits.float() - r_ref.logits.float()).abs().max())
    assert r_our.loss.item() == r_ref.loss.item()
    p_ref, p_our = dict(m_ref.named_parameters()), dict(m_our.named_parameters())
    assert sorted(p_ref) == sorted(p_our)
    checked = 0
    for n in names_our:
        gr, go = p_ref[n + ".weight"].grad, p_our[n + ".weight"].grad
        assert go.dtype == gr.dtype == torch.bfloat16 and go.shape == gr.shape
        d = (go.double() - gr.double()).abs()
        tol = 8 * ulp_of(gr.double(), torch.bfloat16) + 4e-2 * 128 * gr.double().abs().mean()
        assert float((d > tol).float().mean()) < 0.002, (n, float(d.max()))
        rel = float((go.double() - gr.double()).norm() / gr.double().norm())
        assert rel < 2e-2, (n, rel)
        checked += 1
    assert checked == 14
    # non-quantised parameters (embeddings, norms, lm_head) see the same upstream gradients: near-equal
    for n in ["lm_head.weight", "model.norm.weight"]:
        rel = float((p_our[n].grad.double() - p_ref[n].grad.double()).norm() / p_ref[n].grad.double().norm())
        assert rel < 1e-2, (n, rel)


@pytest.mark.parametrize("quant_type", ["int2-asym", "ste-n2f3"])
def test_qlinear_convert_vs_cpu_golden(quant_type, monkeypatch):
    """The same model against the reference run on CPU in the build container (tests/golden/qlinear.npz, CPU semantics):
    bf16 matmuls accumulate in a different order on the two devices, so the bar here is statistical."""
    from bitdistiller_b200 import qlinear as ours
    monkeypatch.setenv("BDQ_SEMANTICS", "cpu")
    z = np.load(GOLDEN + "/qlinear.npz")
    ids = torch.from_numpy(z["ids"]).to(dev)
    model, res = _run_model(ours.convertModelToQuant, quant_type, ids)
    lg, ref_lg = res.logits.float().cpu().double(), torch.from_numpy(z[quant_type + "_logits"]).double()
    assert float((lg - ref_lg).norm() / ref_lg.norm()) < 3e-2
    assert abs(res.loss.item() - float(z[quant_type + "_loss"])) < 2e-2 * float(z[quant_type + "_loss"])
    params = dict(model.named_parameters())
    for n in QL_GRADS:
        g, gr = params[n].grad.float().cpu().double(), torch.from_numpy(z[quant_type + "_g_" + n]).double()
        assert float((g - gr).norm() / gr.norm()) < 8e-2, (n, float((g - gr).norm() / gr.norm()))


def test_qlinear_api(ref):
    from bitdistiller_b200 import qlinear as ours
    for mod in (ours, ref["qlinear"]):
        with pytest.raises(ValueError):
            mod.QLinear(128, 128, quant_type="fp4")
    lin = ours.QLinear(256, 128, bias=True, compute_dtype=torch.bfloat16, quant_type="int2-asym", q_group_size=128, device=dev)
    lin_r = ref["qlinear"].QLinear(256, 128, bias=True, compute_dtype=torch.bfloat16, quant_type="int2-asym", q_group_size=128,
                                   device=dev)
    lin_r.load_state_dict(lin.state_dict())
    x = torch.randn(4, 7, 256, device=dev, dtype=torch.float32)
    y, yr = lin(x), lin_r(x)
    assert y.dtype == torch.float32 and torch.equal(y, yr)


# ---------------------------------------------------------------- a9-a12: packer, GEMV, GEMM against the reference kernels
def _ref_packed(ref, oc, ic, seed, G=128, bias=False):
    """The reference flow of quantizer.py:95-104 on the GPU: pseudo_quantize_tensor -> WQLinear.from_linear."""
    lin = torch.nn.Linear(ic, oc, bias=bias, dtype=torch.float16, device=dev)
    lin.weight.data = make_weight("randn", oc, ic, torch.float16, seed).to(dev)
    w0 = lin.weight.data.clone()
    lin.weight.data, s, z = ref["quantizer"].pseudo_quantize_tensor(lin.weight.data, n_bit=2, zero_point=True,
                                                                    q_group_size=G, get_scale_zp=True)
    ql = ref["qmodule"].WQLinear.from_linear(lin, 2, G, False, s, z)
    return ql, lin, w0


def _ours_packed(lin_w0, oc, ic, G=128, bias=None):
    from bitdistiller_b200.qmodule import WQLinear
    from bitdistiller_b200.quantizer import pseudo_quantize_tensor
    lin = torch.nn.Linear(ic, oc, bias=bias is not None, dtype=torch.float16, device=dev)
    if bias is not None:
        lin.bias.data = bias.clone()
    lin.weight.data, s, z = pseudo_quantize_tensor(lin_w0.clone(), n_bit=2, zero_point=True, q_group_size=G, get_scale_zp=True)
    return WQLinear.from_linear(lin, 2, G, False, s, z), lin


def _tol_check(out, refo, x, dq, what, floor="l1"):
    """rtol 1e-2 (north_star) plus an absolute floor for outputs that are small by cancellation: 2e-3 of the size of a
    random sum of the products -- sum_k |x_k w_k| / sqrt(K) for evenly sized terms (`l1`, the bar of tests/test_packed_gpu.py),
    sqrt(sum_k (x_k w_k)^2) when a few outlier channels carry the sum (`rss`: with 100-1000x channels the l1 form
    underestimates it ~10x).  The batch-1 GEMV quantises x per 128-group to 14-bit block fixed point, so a group that holds a
    1000x channel keeps ~4 bits of its small elements: measured |error| <= 2e-3 on outputs of magnitude ~1."""
    out, refo = out.double(), refo.double()
    if floor == "rss":
        fl = torch.sqrt((x.double() ** 2) @ (dq.double() ** 2).T)
    else:
        fl = (x.double().abs() @ dq.double().abs().T) / np.sqrt(x.shape[1])
    tol = 1e-2 * refo.abs() + 2e-3 * fl + 1e-6
    bad = (out - refo).abs() > tol
    assert not bool(bad.any()), (what, int(bad.sum()), float((out - refo).abs().max()))
    assert float((out - refo).norm() / refo.norm()) < 2e-3, what


@pytest.mark.parametrize("oc,ic", [(256, 512), (512, 4096), (256, 2304)])
def test_packer_vs_reference_packer_on_gpu(ref, oc, ic):
    ql_r, lin_r, w0 = _ref_packed(ref, oc, ic, 3, bias=True)
    ql, lin = _ours_packed(w0, oc, ic, bias=lin_r.bias.data)
    assert torch.equal(lin.weight.data, lin_r.weight.data)
    assert torch.equal(ql.qweight, ql_r.qweight) and torch.equal(ql.qzeros, ql_r.qzeros)
    assert torch.equal(ql.scales, ql_r.scales) and torch.equal(ql.bias, ql_r.bias)
    assert ql.qweight.shape == ql_r.qweight.shape and ql.scales.shape == ql_r.scales.shape


def _x(kind, m, ic, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(m, ic, generator=g)
    if kind == "outlier":   # realistic LLM activations -- per-channel scales plus a few 100-1000x channels
        x = x * torch.exp(torch.randn(ic, generator=g) * 0.5)
        idx = torch.randperm(ic, generator=g)[:6]
        x[:, idx] *= torch.tensor([100.0, 300.0, 1000.0, 100.0, 300.0, 1000.0])
        x = x * 0.02
    return x.to(torch.float16).to(dev)


@pytest.mark.parametrize("oc,ic", [(4096, 4096), (1024, 4096), (4096, 12288), (12288, 4096)])
@pytest.mark.parametrize("kind", ["randn", "outlier"])
def test_gemv_vs_reference_cuda_kernel(ref, oc, ic, kind):
    """a11: bdq_gemv_w2 against the reference's gemv_kernel_g128_2bit (gemv_cuda.cu:202-272) on the SAME packed tensors.
    IC = 11008 is left out: the reference kernel reads activations out of bounds there."""
    from bitdistiller_b200 import awq_inference_engine as E
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq
    w = make_weight("randn", oc, ic, torch.float16, 21).to(dev)
    dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
    qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)    # bit-equal to the reference packer (test above)
    x = _x(kind, 1, ic, 5)
    r = ref["engine"].gemv_forward_cuda(x, qw, qs, qz, 2, 128)
    for flags in (0, 1):
        out = E.gemv_forward_cuda(x, qw, qs, qz, 2, 128, flags=flags)
        assert out.shape == r.shape and out.dtype == r.dtype
        _tol_check(out, r, x, dq, (oc, ic, kind, flags), floor="rss" if kind == "outlier" else "l1")
    # the oracle agrees with the reference kernel too (pins oracle.wq_matmul)
    sub = slice(0, 256)
    o = oracle.wq_matmul(x.cpu(), qw[sub].cpu(), qz[sub].cpu(), qs[sub].cpu(), 128, 2)
    _tol_check(o, r[:, sub].cpu(), x.cpu(), dq[sub].cpu(), "oracle vs reference kernel")


@pytest.mark.parametrize("m,oc,ic", [(64, 256, 512), (200, 1024, 4096), (2048, 4096, 4096), (512, 12288, 4096), (300, 4096, 12288),
                                     (16, 512, 1024)])
@pytest.mark.parametrize("kind", ["randn", "outlier"])
def test_gemm_vs_reference_triton_kernel(ref, m, oc, ic, kind):
    """a12: bdq_gemm_w2 against the reference's Triton quant_matmul_v2 (triton_kernels.py:102-134), called exactly as
    WQLinear.forward calls it (qmodule.py:152-165: transposed contiguous copies of the packed tensors)."""
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200 import triton_kernels as T
    from bitdistiller_b200.qmodule import pack_wq
    w = make_weight("randn", oc, ic, torch.float16, 5).to(dev)
    dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
    qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
    x = _x(kind, m, ic, m)
    kw = dict(M=m, N=oc, K=ic, pack_num=16, group_size=128, w_bit=2, offset=3)
    r = ref["triton_kernels"].quant_matmul_v2(x, qw.T.contiguous(), qz.T.contiguous(), qs.T.contiguous(), **kw)
    torch.cuda.synchronize()
    out = T.gemm_w2(x, qw, qs, qz, 128)
    _tol_check(out, r, x, dq, (m, oc, ic, kind, "gemm_w2"))
    out2 = T.quant_matmul_v2(x, qw.T.contiguous(), qz.T.contiguous(), qs.T.contiguous(), **kw)   # the drop-in signature
    assert torch.equal(out2, out)
    sub = slice(0, 64)
    o = oracle.wq_matmul(x[:8].cpu(), qw[sub].cpu(), qz[sub].cpu(), qs[sub].cpu(), 128, 2, dequant_fp16=True)
    _tol_check(o, r[:8, sub].cpu(), x[:8].cpu(), dq[sub].cpu(), "oracle vs reference Triton kernel")


@pytest.mark.parametrize("shape", [(1, 4096), (1, 1, 4096), (2, 64, 4096), (77, 4096)])
def test_wqlinear_forward_vs_reference_module(ref, shape):
    """a10: WQLinear.forward (qmodule.py:148-173) module against module, bias included; rows == 1 takes the reference's
    CUDA GEMV, more rows its Triton GEMM (USE_TRITON is True in this image).  x.shape[-2] == rows except for batch > 1,
    where the reference passes the wrong M: compare the rows it does compute."""
    oc, ic = 1024, shape[-1]
    ql_r, lin_r, w0 = _ref_packed(ref, oc, ic, 9, bias=True)
    ql, lin = _ours_packed(w0, oc, ic, bias=lin_r.bias.data)
    x = _x("randn", int(np.prod(shape[:-1])), ic, 3).reshape(shape)
    y = ql(x)
    assert y.shape == shape[:-1] + (oc,) and y.dtype == torch.float16
    if len(shape) == 3 and shape[0] > 1:
        yr = torch.stack([ql_r(x[i:i + 1])[0] for i in range(shape[0])])   # per batch row: M == x.shape[-2] is then right
    else:
        yr = ql_r(x)
    _tol_check(y.reshape(-1, oc), yr.reshape(-1, oc), x.reshape(-1, ic), lin.weight.data, shape)


# ---------------------------------------------------------------- a7: the clip search against the reference on the GPU
@pytest.mark.parametrize("qt", ["int", "nf3"])
def test_autoclip_vs_reference_on_gpu(qt):
    """auto_2clip_layer (autoclip.py:13-83) run by the reference on the B200 in bf16 against ours, judged under the
    oracle's exact error (see tests/test_autoclip_gpu.py for why index equality with a bf16 run is not defined)."""
    import importlib
    load()
    try:
        ra = importlib.import_module("autoclip")   # the reference's module (pulls datasets / transformers)
    except Exception as e:  # pragma: no cover
        pytest.skip(f"reference autoclip.py does not import on this box: {type(e).__name__}: {e}")
    from bitdistiller_b200.autoclip import auto_2clip_layer
    g = torch.Generator().manual_seed(23)
    G, co, ci = 128, 64, 512
    w = (torch.randn(co, ci, generator=g) * 0.02).to(torch.bfloat16)
    x = (torch.randn(1024, ci, generator=g) * torch.exp(torch.randn(ci, generator=g) * 0.5)).to(torch.bfloat16)
    q_config = {"zero_point": True, "q_group_size": G, "quant_type": qt}
    rmax, rmin = ra.auto_2clip_layer(w.to(dev).clone(), x.to(dev).clone(), n_bit=2, q_config=q_config)
    bmax, bmin, idx, err = auto_2clip_layer(w.to(dev), x.to(dev), 2, q_config, return_info=True)
    r = oracle.auto_2clip(w, x[0::2], G, 2, qt, sem=1, want_all_err=True)
    wg = w.reshape(co, -1, G)
    fac = torch.tensor([float(np.float32(1 - i / 20)) for i in range(10)])
    cmax = (wg.amax(-1).float()[..., None] * fac).to(torch.bfloat16).float()
    cmin = (wg.amin(-1).float()[..., None] * fac).to(torch.bfloat16).float()
    hp, hn = cmax == rmax.cpu().float().reshape(co, -1, 1), cmin == rmin.cpu().float().reshape(co, -1, 1)
    assert bool(hp.any(-1).all() and hn.any(-1).all())
    ridx = hp.float().argmax(-1) * 10 + hn.float().argmax(-1)
    ii = idx.cpu().long().clamp(min=0)
    e_ref = r["all_err"].gather(-1, ridx[..., None]).squeeze(-1)
    e_our = r["all_err"].gather(-1, ii[..., None]).squeeze(-1)
    assert bool((e_our <= e_ref * (1 + 1e-3) + 1e-30).all()), float((e_our / e_ref).max())
    assert bool((e_ref <= e_our * 1.05 + 1e-30).all()), float((e_ref / e_our).max())
    assert float((ii == r["idx"].long()).float().mean()) >= 0.985
    print(qt, "agreement with the reference's bf16 choice on the GPU:", float((ii == ridx).float().mean()))
