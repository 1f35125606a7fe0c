"""Decode GEMV micro-benchmark (BASELINE config 5a): Qwen3-8B packed int2-g128, batch 1, all 36 x 7 linears with
distinct weight memory (1.86 GB per token >> 126 MB L2 => cache-cold).  Modes: eager stream launches, PDL launches,
CUDA graph, CUDA graph + PDL.  Prints tok/s and the fraction of the measured HBM peak (0.26758 B/weight)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
from bitdistiller_b200 import quantizer as Q  # noqa: E402
from bitdistiller_b200.qmodule import pack_wq  # noqa: E402

QWEN3_8B = [("q_proj", 4096, 4096), ("k_proj", 1024, 4096), ("v_proj", 1024, 4096), ("o_proj", 4096, 4096),
            ("gate_proj", 12288, 4096), ("up_proj", 12288, 4096), ("down_proj", 4096, 12288)]
LAYERS = 36


def build(dev, layers=LAYERS, batch=1, chain=False):
    """chain=True: a truly dependent decode chain -- every GEMV reads the activation vector the previous launch just
    wrote (q_proj's output feeds o_proj, o_proj's output feeds gate/up, gate_proj's output feeds down_proj, down_proj's
    output feeds the next layer's q/k/v; the attention / SiLU math in between is not part of this path), weights scaled
    so the chain keeps unit variance.  chain=False: every launch has its own static input."""
    proto = {}
    for name, oc, ic in QWEN3_8B:
        if (oc, ic) in proto:
            continue
        g = torch.Generator(device=dev).manual_seed(oc * 7 + ic)
        w = (torch.randn(oc, ic, device=dev, generator=g) * (ic ** -0.5 if chain else 0.02)).half()
        dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
        proto[(oc, ic)] = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
    ops = []
    for li in range(layers):
        for name, oc, ic in QWEN3_8B:
            qw, qz, qs = (t.clone() for t in proto[(oc, ic)])
            x = torch.randn(batch, ic, device=dev).half()
            out = torch.empty(batch, oc, device=dev, dtype=torch.float16)
            ops.append((x, qw, qs, qz, out, batch, ic, oc))
    if chain:
        assert batch == 1
        prev_down = None
        for li in range(layers):
            o = {n[0]: ops[li * 7 + k] for k, n in enumerate(QWEN3_8B)}
            src = {"q_proj": prev_down, "k_proj": prev_down, "v_proj": prev_down, "o_proj": o["q_proj"][4],
                   "gate_proj": o["o_proj"][4], "up_proj": o["o_proj"][4], "down_proj": o["gate_proj"][4]}
            for k, n in enumerate(QWEN3_8B):
                if src[n[0]] is not None:
                    t = ops[li * 7 + k]
                    ops[li * 7 + k] = (src[n[0]],) + t[1:]
            prev_down = o["down_proj"][4]
    return ops


def bytes_per_token(ops):
    tot = 0
    for x, qw, qs, qz, out, b, ic, oc in ops:
        tot += oc * ic * 0.25 + oc * (ic // 128) * 2 + oc * (ic // 128) * 0.25  # codes + fp16 scales + 2-bit zeros
    return tot


def groups_of(ops, fused):
    """Launch groups: [(first op index, n matrices)]."""
    out, i = [], 0
    while i < len(ops):
        n = 1
        if fused and ops[i][5] == 1:
            name = QWEN3_8B[i % 7][0]
            n = 3 if name == "q_proj" else (2 if name == "gate_proj" else 1)
        out.append((i, n))
        i += n
    return out


def run(ops, flags, st, fused=False, prefetch=False):
    """One token's worth of decode GEMVs.  fused=True: q/k/v and gate/up go through one bdq_gemv_w2_multi launch each
    (4 launches per layer instead of 7).  prefetch=True: every launch carries the packed weights of the next launch
    as its L2 prefetch hint."""
    import ctypes
    lib = L.lib()
    gl = groups_of(ops, fused)
    for gi, (i, n) in enumerate(gl):
        x, qw, qs, qz, out, b, ic, oc = ops[i]
        if b != 1 or (n == 1 and not prefetch):
            rc = lib.bdq_gemv_w2(L.ptr(x), L.ptr(qw), L.ptr(qs), L.ptr(qz), L.ptr(out), L.i64(b), L.i64(ic), L.i64(oc),
                                 128, flags, st)
        else:
            grp = ops[i:i + n]
            arr = lambda k: (ctypes.c_void_p * n)(*[o[k].data_ptr() for o in grp])
            ocs = (ctypes.c_int64 * n)(*[o[7] for o in grp])
            npf, pfp, pfb = 0, None, None
            if prefetch and gi + 1 < len(gl):
                j, nn = gl[gi + 1]
                nxt = [o[1] for o in ops[j:j + nn]]
                npf = len(nxt)
                pfp = (ctypes.c_void_p * npf)(*[t.data_ptr() for t in nxt])
                pfb = (ctypes.c_int64 * npf)(*[t.numel() * 4 for t in nxt])
            rc = lib.bdq_gemv_w2_multi(L.ptr(x), n, arr(1), arr(2), arr(3), arr(4), ocs, L.i64(ic), 128, flags, npf, pfp,
                                       pfb, st)
        L.check(rc, "bdq_gemv_w2")


def time_fn(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2] * 1e-3


def chain_of(ops):
    """The same dependent ops as ONE persistent launch (DecodeChain): phases = the fused launch groups."""
    from bitdistiller_b200.awq_inference_engine import DecodeChain
    phases = []
    for gi, (i, n) in enumerate(groups_of(ops, True)):
        grp = ops[i:i + n]
        phases.append({"x": grp[0][0], "mats": [(o[1], o[2], o[3], o[4]) for o in grp], "nowait": gi == 0})
    return DecodeChain(phases)


def chain_time(ops, lookahead=2, iters=20):
    ch = chain_of(ops)
    return time_fn(lambda: ch.run(lookahead), iters=iters)


def graph_time(ops, flags, fused, dev, prefetch=False):
    s = torch.cuda.Stream(dev)
    with torch.cuda.stream(s):
        run(ops, flags, ctypes_stream(s), fused, prefetch)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=s):
            run(ops, flags, ctypes_stream(s), fused, prefetch)
    return time_fn(gr.replay)


def main():
    dev = torch.device("cuda:0")
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    chain = batch == 1 and not (len(sys.argv) > 2 and sys.argv[2] == "static")
    ops = build(dev, batch=batch, chain=chain)
    print(f"inputs: {'dependent chain (each launch reads the previous launch output)' if chain else 'static per-launch inputs'}")
    peak = 6548.5
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    nbytes = bytes_per_token(ops)
    res = {}
    PDL, C1, RING = 1, 2, 4
    variants = [("ring", RING), ("stream", 0), ("stream tma", 16)] if batch == 1 else [("ring", 0)]
    for vname, vf in variants:
        for fused in ((False, True) if batch == 1 else (False,)):
            for pname, pf in (("graph", 0), ("graph+pdl", PDL)):
                for pre in ((False, True) if (batch == 1 and vf != RING) else (False,)):
                    key = f"{vname:14s} {'fused(4/layer)' if fused else 'split(7/layer)'} {pname:9s} {'+L2hint' if pre else ''}"
                    res[key] = graph_time(ops, vf | pf, fused, dev, pre)
    if batch == 1:
        res["persistent chain (1 launch/token)"] = chain_time(ops, 0)
    for k, t in res.items():
        print(f"batch {batch} {k:50s}: {t*1e6:8.1f} us/token  {1/t:8.1f} tok/s  {nbytes/t/1e9:7.1f} GB/s  "
              f"{nbytes/t/1e9/peak*100:5.1f}% of measured HBM peak ({nbytes/1e9:.3f} GB/token)")
    best = min(res, key=res.get)
    print(json.dumps({"best": best.strip(), "us_per_token": res[best] * 1e6, "tok_s": 1 / res[best],
                      "frac_hbm_peak": nbytes / res[best] / 1e9 / peak, "bytes_per_token": nbytes, "peak_gbs": peak}))
    # per-shape timing inside a graph of 36 launches over the 36 distinct layer copies (cold), stream kernel
    for name, oc, ic in QWEN3_8B:
        sel = [o for o in ops if o[6] == ic and o[7] == oc][:36]
        b = oc * ic * 0.26758
        for pname, pf in (("graph", 0), ("graph+pdl", PDL)):
            t = graph_time(sel, pf, False, dev) / len(sel)
            print(f"  {name:10s} [{oc}x{ic}] {pname:9s} {t*1e6:7.2f} us/launch  {b/t/1e9:7.1f} GB/s  {b/t/1e9/peak*100:5.1f}%")


def ctypes_stream(s):
    import ctypes
    return ctypes.c_void_p(s.cuda_stream)


if __name__ == "__main__":
    main()
