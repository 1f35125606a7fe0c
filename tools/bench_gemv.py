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


def build(dev, layers=LAYERS, batch=1):
    proto = {}
    for name, oc, ic in QWEN3_8B:
        if (oc, ic) in proto:
            continue
        g = torch.Generator(device=dev).manual_seed(oc * 7 + ic)
        w = (torch.randn(oc, ic, device=dev, generator=g) * 0.02).half()
        dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
        proto[(oc, ic)] = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
    ops = []
    for li in range(layers):
        for name, oc, ic in QWEN3_8B:
            qw, qz, qs = (t.clone() for t in proto[(oc, ic)])
            x = torch.randn(batch, ic, device=dev).half()
            out = torch.empty(batch, oc, device=dev, dtype=torch.float16)
            ops.append((x, qw, qs, qz, out, batch, ic, oc))
    return ops


def bytes_per_token(ops):
    tot = 0
    for x, qw, qs, qz, out, b, ic, oc in ops:
        tot += oc * ic * 0.25 + oc * (ic // 128) * 2 + oc * (ic // 128) * 0.25  # codes + fp16 scales + 2-bit zeros
    return tot


def run(ops, flags, st):
    lib = L.lib()
    for x, qw, qs, qz, out, b, ic, oc in ops:
        rc = lib.bdq_gemv_w2(L.ptr(x), L.ptr(qw), L.ptr(qs), L.ptr(qz), L.ptr(out), L.i64(b), L.i64(ic), L.i64(oc), 128,
                             flags, st)
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


def main():
    dev = torch.device("cuda:0")
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    ops = build(dev, batch=batch)
    peak = 6548.5
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    nbytes = bytes_per_token(ops)
    st = L.stream_ptr(dev)
    res = {}
    for name, flags in [("eager", 0), ("eager+pdl", 1)]:
        t = time_fn(lambda: run(ops, flags, st))
        res[name] = t
    for name, flags in [("graph", 0), ("graph+pdl", 1)]:
        s = torch.cuda.Stream(dev)
        with torch.cuda.stream(s):
            run(ops, flags, L.stream_ptr(dev))
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                run(ops, flags, ctypes_stream(s))
        t = time_fn(gr.replay)
        res[name] = t
    for k, t in res.items():
        print(f"batch {batch} {k:10s}: {t*1e6:8.1f} us/token  {1/t:8.1f} tok/s  {nbytes/t/1e9:7.1f} GB/s  "
              f"{nbytes/t/1e9/peak*100:5.1f}% of measured HBM peak ({nbytes/1e9:.3f} GB/token)")
    # per-shape isolated timing (rotating over the 36 layer copies => cold)
    for name, oc, ic in QWEN3_8B:
        sel = [o for o in ops if o[6] == ic and o[7] == oc]
        b = sel[0][7] * sel[0][6] * 0.26758
        for flags in (0,):
            t = time_fn(lambda: run(sel, flags, st)) / len(sel)
            print(f"  {name:10s} [{oc}x{ic}] {t*1e6:7.2f} us/launch  {b/t/1e9:7.1f} GB/s  {b/t/1e9/peak*100:5.1f}%")


def ctypes_stream(s):
    import ctypes
    return ctypes.c_void_p(s.cuda_stream)


if __name__ == "__main__":
    main()
