"""Packed NF3 decode GEMV timing (0.40625 B/weight at group 128), L2-cold: each launch reads a different copy of the packed
matrix, >= 400 MB of copies per shape, the launches of one pass captured in a CUDA graph.  BDQ_NF3_GEMV_STREAM=0 times the pair-lookup kernel (the round-1 path) instead of
the PRMT-lookup / INT8-MMA streaming kernel.     python tools/bench_nf3_gemv.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import nf3_packed as P  # noqa: E402
from tools.bench_gemv import time_fn  # noqa: E402

SHAPES = [(4096, 4096), (12288, 4096), (4096, 12288), (32000, 4096), (151936, 4096)]


def run(dev, peak, shapes=SHAPES):
    res = []
    for oc, ic in shapes:
        nb = oc * ic * 0.40625
        ncopy = max(2, int(400e6 // nb) + 1)
        mats = []
        for _ in range(ncopy):
            qw = torch.randint(-2**31, 2**31 - 1, (oc, ic // 16), device=dev, dtype=torch.int32)
            qh = torch.randint(-2**31, 2**31 - 1, (oc, ic // 32), device=dev, dtype=torch.int32)
            sc = (torch.rand(oc, 2 * ic // 128, device=dev) * 0.01 + 0.001).half()
            mats.append((qw, qh, sc))
        x = torch.randn(1, ic, device=dev).half()

        def once():
            for qw, qh, sc in mats:
                P.gemv_nf3(x, qw, qh, sc, 128)
        once()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()     # the small shapes are shorter than an eager launch from Python
        with torch.cuda.graph(graph):
            once()
        t = time_fn(graph.replay) / ncopy
        res.append({"shape": [oc, ic], "mb": nb / 1e6, "us": t * 1e6, "gbs": nb / t / 1e9, "frac_of_hbm_peak": nb / t / 1e9 / peak})
        del mats
    return res


if __name__ == "__main__":
    pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
    peak = json.load(open(pk)).get("hbm_gbs", 6462.4) if os.path.exists(pk) else 6462.4
    shapes = SHAPES[-2:] if len(sys.argv) > 1 and sys.argv[1] == "big" else SHAPES
    for r in run(torch.device("cuda:0"), peak, shapes):
        print(f"nf3 gemv [{r['shape'][0]:6d} x {r['shape'][1]:5d}] {r['mb']:7.1f} MB  {r['us']:8.2f} us/launch  {r['gbs']:7.1f} GB/s  "
              f"{r['frac_of_hbm_peak']*100:5.1f}% of HBM peak")
