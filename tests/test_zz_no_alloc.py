"""bdq_gemv_w2_chain_static (include/bdq.h): the decode chain with its scratch in a file-scope __device__ array -- the entry
point for callers that cannot cudaMalloc.  No counterpart in the reference;
the oracle here is the existing entry point pair bdq_gemv_w2_chain_prepare / _run on the same tensors, compared bit for
bit (the tolerance test_decode_chain_equals_launch_sequence uses: same kernel, same arithmetic).

Named test_zz_* so that pytest collects it last: `pytest -x` must not lose other tests to a failure here.
It passes alone in a fresh process and after the other tests (it keeps no state of its own; the static array is rewritten
at every call)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
dev = torch.device("cuda:0")


def _packed(oc, ic, seed):
    from bitdistiller_b200 import quantizer as Q
    from bitdistiller_b200.qmodule import pack_wq
    g = torch.Generator(device=dev).manual_seed(seed)
    w = (torch.randn(oc, ic, device=dev, generator=g) * ic ** -0.5).half()
    dq, sc, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
    qw, qz, qs = pack_wq(dq, sc.view(oc, -1), z.view(oc, -1), 2, 128)
    return qw, qs, qz, torch.zeros(1, oc, dtype=torch.float16, device=dev)


def _chain(hidden, kv, inter, layers, silu, seed):
    """q/k/v <- x, o <- q, gate/up <- o, down <- gate (or silu(gate) * up), next layer <- down: every phase depends on the
    one before it, so a stale counter or plan in the shared scratch shows up as a wrong output."""
    x = torch.randn(1, hidden, generator=torch.Generator().manual_seed(seed)).half().to(dev)
    phases = []
    for li in range(layers):
        s = 100 * seed + 10 * li
        q, k, v, o = _packed(hidden, hidden, s + 1), _packed(kv, hidden, s + 2), _packed(kv, hidden, s + 3), _packed(hidden, hidden, s + 4)
        gt, up, dn = _packed(inter, hidden, s + 5), _packed(inter, hidden, s + 6), _packed(hidden, inter, s + 7)
        phases.append({"x": x, "mats": [q, k, v], "nowait": li == 0})
        phases.append({"x": q[3], "mats": [o]})
        phases.append({"x": o[3], "mats": [gt, (up[0], up[1], up[2], None)], "silu": True} if silu else {"x": o[3], "mats": [gt, up]})
        phases.append({"x": gt[3], "mats": [dn]})
        x = dn[3]
    return phases


def _phase_array(phases):
    from bitdistiller_b200 import awq_inference_engine as E
    arr = (E._Phase * len(phases))()
    for a, ph in zip(arr, phases):
        x, mats = ph["x"], ph["mats"]
        a.x, a.nmat, a.group, a.IC = x.data_ptr(), len(mats), 128, x.shape[-1]
        a.flags = (E.GEMV_SILU_MUL if ph.get("silu") else 0) | (E.GEMV_CHAIN_NOWAIT if ph.get("nowait") else 0)
        for i, (qw, sc, qz, out) in enumerate(mats):
            a.qweight[i], a.scales[i], a.qzeros[i] = qw.data_ptr(), sc.data_ptr(), qz.data_ptr()
            a.out[i] = 0 if out is None else out.data_ptr()
            a.OC[i] = qw.shape[0]
    return arr


def _outs(phases):
    return [m[3] for ph in phases for m in ph["mats"] if m[3] is not None]


def _run_static(arr, n):
    from bitdistiller_b200 import _lib as L
    with torch.cuda.device(dev):
        return L.lib().bdq_gemv_w2_chain_static(arr, ctypes.c_int(n), ctypes.c_int(0), L.stream_ptr(dev))


def test_chain_static_matches_workspace_entry_point():
    from bitdistiller_b200 import _lib as L
    from bitdistiller_b200 import awq_inference_engine as E
    chains = [_chain(1024, 256, 2816, 2, False, seed=1),     # 8 phases
              _chain(512, 128, 1408, 4, True, seed=2),       # 16 phases, silu(gate) * up fused
              _chain(4096, 1024, 12288, 1, False, seed=3)]   # 4 phases at the Qwen3-8B shapes (plan shorter than the one before)
    refs, arrs = [], []
    for phases in chains:    # the existing entry points (caller-provided workspace) are the oracle
        E.DecodeChain(phases).run()
        torch.cuda.synchronize()
        refs.append([o.clone() for o in _outs(phases)])
        assert all(bool(r.isfinite().all()) and float(r.abs().max()) > 0 for r in refs[-1])
        arrs.append(_phase_array(phases))
    # calls in a row on the one shared array: different plans, different phase counts, then the first one again
    for rep, ci in enumerate([0, 1, 2, 0, 0]):
        phases = chains[ci]
        for o in _outs(phases):
            o.zero_()
        torch.cuda.synchronize()
        alloc0, free0 = torch.cuda.memory_allocated(dev), torch.cuda.mem_get_info(dev)[0]
        L.check(_run_static(arrs[ci], len(phases)), "bdq_gemv_w2_chain_static")
        torch.cuda.synchronize()
        if rep > 0:   # (the first call may load the kernel image lazily)
            assert torch.cuda.memory_allocated(dev) == alloc0 and torch.cuda.mem_get_info(dev)[0] >= free0, "the call allocated device memory"
        for i, (a, b) in enumerate(zip(_outs(phases), refs[ci])):
            assert torch.equal(a, b), (rep, ci, i, float((a.float() - b.float()).abs().max()))


def test_chain_static_rejects_oversized_and_bad_plans():
    from bitdistiller_b200 import _lib as L
    phases = _chain(512, 128, 1408, 1, False, seed=4)
    arr1 = _phase_array(phases)
    many = (type(arr1[0]) * 65)()
    for i in range(65):
        ctypes.memmove(ctypes.byref(many[i]), ctypes.byref(arr1[i % 4]), ctypes.sizeof(arr1[0]))
    rc = _run_static(many, 65)
    assert rc != 0
    msg = L.lib().bdq_last_error().decode()
    assert "BDQ_GEMV_CHAIN_STATIC_MAX_PHASES" in msg and "65" in msg, msg
    with pytest.raises(L.BdqError):
        L.check(_run_static(arr1, 0), "bdq_gemv_w2_chain_static")
    bad = _phase_array(phases)
    bad[1].group = 64
    with pytest.raises(L.BdqError, match="group 128"):
        L.check(_run_static(bad, 4), "bdq_gemv_w2_chain_static")
    # the array is still usable after the rejected calls
    from bitdistiller_b200 import awq_inference_engine as E
    E.DecodeChain(phases).run()
    torch.cuda.synchronize()
    E_ref = [o.clone() for o in _outs(phases)]
    for o in _outs(phases):
        o.zero_()
    L.check(_run_static(arr1, 4), "bdq_gemv_w2_chain_static")
    torch.cuda.synchronize()
    assert all(torch.equal(a, b) for a, b in zip(_outs(phases), E_ref))
