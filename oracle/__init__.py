"""CPU oracle (ctypes wrapper over oracle/bdq_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, ``__graft_entry__.smoke()`` and the cpu_baseline /
``--impl reference`` legs of bench.py.  The product package ``bitdistiller_b200`` never imports it.

Each function restates one reference function (file:line cited in bdq_oracle.c) and is pinned against
outputs of the unmodified reference stored in tests/golden/*.npz (tests/test_oracle_golden.py).
"""
import ctypes
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libbdq_oracle.so")

_DT = {torch.float32: 0, torch.bfloat16: 1, torch.float16: 2}


def build(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "bdq_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-B"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
    return _lib


def _f32(t):
    """torch tensor (any float dtype) -> contiguous float32 numpy holding the same values."""
    return np.ascontiguousarray(t.detach().cpu().to(torch.float32).numpy())


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def int_fakequant(w, G, bits, round_mode=0, sem=0, clip_min=None, clip_max=None):
    """-> dict(dq, scales, zeros, codes, mask); dq/scales/zeros in w.dtype."""
    dt = _DT[w.dtype]
    x = _f32(w)
    n = x.size
    if G <= 0:
        G = w.shape[-1]
    dq = np.empty(n, np.float32)
    s = np.empty(n // G, np.float32)
    z = np.empty(n // G, np.float32)
    codes = np.empty(n, np.uint8)
    mask = np.empty(n, np.uint8)
    cmin = None if clip_min is None else _f32(clip_min).reshape(-1)
    cmax = None if clip_max is None else _f32(clip_max).reshape(-1)
    lib().orc_int_fakequant(_p(x), ctypes.c_int64(n), G, bits, dt, round_mode, sem, _p(cmin), _p(cmax),
                            _p(dq), _p(s), _p(z), _p(codes), _p(mask))
    rows = x.reshape(-1, w.shape[-1]).shape[0]
    return dict(dq=torch.from_numpy(dq).reshape(w.shape).to(w.dtype),
                scales=torch.from_numpy(s).reshape(rows, -1).to(w.dtype),
                zeros=torch.from_numpy(z).reshape(rows, -1).to(w.dtype),
                codes=torch.from_numpy(codes).reshape(w.shape),
                mask=torch.from_numpy(mask).reshape(w.shape))


def int_ste_bwd(w, gy, G, bits, sem=0):
    """-> float64 gradient tensor of w's shape."""
    dt = _DT[w.dtype]
    x, g = _f32(w), _f32(gy)
    if G <= 0:
        G = w.shape[-1]
    gw = np.empty(x.size, np.float64)
    lib().orc_int_ste_bwd(_p(x), _p(g), ctypes.c_int64(x.size), G, bits, dt, sem, _p(gw))
    return torch.from_numpy(gw).reshape(w.shape)


def nf3_fakequant(w, G, sem=0, clip_min=None, clip_max=None):
    dt = _DT[w.dtype]
    x = _f32(w)
    n = x.size
    if G <= 0:
        G = w.shape[-1]
    dq = np.empty(n, np.float32)
    sp = np.empty(n // G, np.float32)
    sn = np.empty(n // G, np.float32)
    codes = np.empty(n, np.uint8)
    cmin = None if clip_min is None else _f32(clip_min).reshape(-1)
    cmax = None if clip_max is None else _f32(clip_max).reshape(-1)
    lib().orc_nf3_fakequant(_p(x), ctypes.c_int64(n), G, dt, sem, _p(cmin), _p(cmax), _p(dq), _p(sp), _p(sn), _p(codes))
    return dict(dq=torch.from_numpy(dq).reshape(w.shape).to(w.dtype),
                scale_pos=torch.from_numpy(sp).to(w.dtype), scale_neg=torch.from_numpy(sn).to(w.dtype),
                codes=torch.from_numpy(codes).reshape(w.shape))


def nf3_ste_bwd(w, gy, G, sem=0):
    dt = _DT[w.dtype]
    x, g = _f32(w), _f32(gy)
    if G <= 0:
        G = w.shape[-1]
    gw = np.empty(x.size, np.float64)
    lib().orc_nf3_ste_bwd(_p(x), _p(g), ctypes.c_int64(x.size), G, dt, sem, _p(gw))
    return torch.from_numpy(gw).reshape(w.shape)


def apply_clip(w, clip_min, clip_max):
    G = w.numel() // clip_min.numel()
    x = _f32(w).copy()
    lib().orc_apply_clip(_p(x), ctypes.c_int64(x.size), G, _p(_f32(clip_min).reshape(-1)), _p(_f32(clip_max).reshape(-1)))
    return torch.from_numpy(x).reshape(w.shape).to(w.dtype)


def auto_2clip(w, x_sampled, G, bits=2, quant_type="int", n_grid=20, max_shrink=0.5, sem=0, want_all_err=False):
    """w [co, ci]; x_sampled [n_tok, ci] already sub-sampled. -> dict(best_max, best_min [co, n_g, 1], idx, err[, all_err])."""
    dt = _DT[w.dtype]
    co, ci = w.shape
    if G <= 0:
        G = ci
    ng = ci // G
    nc = int(max_shrink * n_grid)
    ww, xx = _f32(w), _f32(x_sampled.reshape(-1, ci))
    bmax = np.empty(co * ng, np.float32)
    bmin = np.empty(co * ng, np.float32)
    bidx = np.empty(co * ng, np.int32)
    berr = np.empty(co * ng, np.float64)
    aerr = np.empty(co * ng * nc * nc, np.float64) if want_all_err else None
    lib().orc_auto_2clip(_p(ww), _p(xx), ctypes.c_int64(co), ctypes.c_int64(ci), ctypes.c_int64(xx.shape[0]), G, bits,
                         dt, sem, 0 if quant_type == "int" else 1, n_grid, ctypes.c_double(max_shrink),
                         _p(bmax), _p(bmin), _p(bidx), _p(berr), _p(aerr))
    out = dict(best_max=torch.from_numpy(bmax).reshape(co, ng, 1).to(w.dtype),
               best_min=torch.from_numpy(bmin).reshape(co, ng, 1).to(w.dtype),
               idx=torch.from_numpy(bidx).reshape(co, ng), err=torch.from_numpy(berr).reshape(co, ng))
    if want_all_err:
        out["all_err"] = torch.from_numpy(aerr).reshape(co, ng, nc * nc)
    return out


def zeros_width(in_features, G, pack):
    return lib().orc_zeros_width(ctypes.c_int64(in_features), G, pack)


def pack(w_dq, scales, zeros, G, w_bit):
    """-> (qweight int32 [rows, cols/pack], qzeros int32 [rows, zw], qscales fp16 [rows, zw*pack])."""
    dt = _DT[w_dq.dtype]
    rows, cols = w_dq.shape
    if G <= 0:
        G = cols
    pk = 32 // w_bit
    zw = zeros_width(cols, G, pk)
    qw = np.empty((rows, cols // pk), np.int32)
    qz = np.empty((rows, zw), np.int32)
    qs = np.empty((rows, zw * pk), np.float32)
    lib().orc_pack(_p(_f32(w_dq)), _p(_f32(scales)), _p(_f32(zeros)), ctypes.c_int64(rows), ctypes.c_int64(cols), G,
                   w_bit, dt, _p(qw), _p(qz), _p(qs))
    return torch.from_numpy(qw), torch.from_numpy(qz), torch.from_numpy(qs).to(torch.float16)


def wq_matmul(x, qweight, qzeros, qscales, G, w_bit, dequant_fp16=False):
    """x [B, IC] fp16 -> out [B, OC] fp16 (double accumulation)."""
    B, IC = x.shape
    OC = qweight.shape[0]
    out = np.empty((B, OC), np.float32)
    qw = np.ascontiguousarray(qweight.cpu().numpy().astype(np.int32))
    qz = np.ascontiguousarray(qzeros.cpu().numpy().astype(np.int32))
    lib().orc_wq_matmul(_p(_f32(x)), _p(qw), _p(qz), _p(_f32(qscales)), ctypes.c_int64(B), ctypes.c_int64(IC),
                        ctypes.c_int64(OC), G, w_bit, int(dequant_fp16), _p(out))
    return torch.from_numpy(out).to(torch.float16)


def gptq_export_w2(w, G):
    """w [OC, IC] -> (qweight int32 [IC/16, OC], qzeros int32 [IC/G, OC/16], scales bf16 [IC/G, OC])."""
    dt = _DT[w.dtype]
    oc, ic = w.shape
    qw = np.empty((ic // 16, oc), np.int32)
    qz = np.empty((ic // G, oc // 16), np.int32)
    sc = np.empty((ic // G, oc), np.float32)
    lib().orc_gptq_export_w2(_p(_f32(w)), ctypes.c_int64(oc), ctypes.c_int64(ic), G, dt, _p(qw), _p(qz), _p(sc))
    return torch.from_numpy(qw), torch.from_numpy(qz), torch.from_numpy(sc).to(torch.bfloat16)
