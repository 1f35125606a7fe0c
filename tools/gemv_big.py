"""Streaming roofline of the batch-1 int2 GEMV kernel itself: one launch over matrices large enough that the fixed
per-launch latency (dependency resolution, x staging, reduction) is small against the streaming time.  Each timed graph
replays launches over distinct weight copies (> L2 in total).  usage: python tools/gemv_big.py [flags]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L  # noqa: E402
from tools.bench_gemv import time_fn, ctypes_stream  # noqa: E402


def run(dev, peak, flags=1, shapes=((12288, 4096), (49152, 4096), (151936, 4096), (65536, 12288))):
    lib = L.lib()
    res = []
    for oc, ic in shapes:
        nbytes = oc * ic * 0.26758
        ncopy = max(2, int(400e6 // nbytes) + 1)
        g = torch.Generator(device=dev).manual_seed(oc + ic)
        mats = []
        for _ in range(ncopy):
            qw = torch.randint(-2**31, 2**31 - 1, (oc, ic // 16), device=dev, dtype=torch.int32, generator=g)
            zw = max(1, (ic // 128 + 15) // 16)
            qz = torch.randint(-2**31, 2**31 - 1, (oc, zw), device=dev, dtype=torch.int32, generator=g)
            sc = (torch.rand(oc, zw * 16, device=dev, generator=g) * 0.01 + 0.001).half()
            mats.append((qw, sc, qz))
        x = torch.randn(1, ic, device=dev).half()
        out = torch.empty(1, oc, device=dev, dtype=torch.float16)
        s = torch.cuda.Stream(dev)

        def launches():
            for qw, sc, qz in mats:
                L.check(lib.bdq_gemv_w2(L.ptr(x), L.ptr(qw), L.ptr(sc), L.ptr(qz), L.ptr(out), L.i64(1), L.i64(ic), L.i64(oc),
                                        128, flags, ctypes_stream(s)), "bdq_gemv_w2")
        with torch.cuda.stream(s):
            launches()
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                launches()
        t = time_fn(gr.replay) / ncopy
        res.append({"OC": oc, "IC": ic, "MB": nbytes / 1e6, "us_per_launch": t * 1e6, "gbs": nbytes / t / 1e9,
                    "frac_of_hbm_peak": nbytes / t / 1e9 / peak, "copies": ncopy})
        del mats
    return res


def main():
    dev = torch.device("cuda:0")
    flags = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
    peak = json.load(open(pk)).get("hbm_gbs", 6553.3) if os.path.exists(pk) else 6553.3
    for r in run(dev, peak, flags):
        print(f"flags {flags:2d} [{r['OC']:6d} x {r['IC']:5d}] {r['MB']:7.1f} MB  {r['us_per_launch']:8.2f} us/launch  "
              f"{r['gbs']:7.1f} GB/s  {r['frac_of_hbm_peak']*100:5.1f}% of HBM peak  ({r['copies']} copies per graph)")


if __name__ == "__main__":
    main()
