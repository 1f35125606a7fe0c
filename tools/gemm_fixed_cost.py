"""Fixed cost of one bdq_gemm_w2 launch: time against the number of k blocks at a fixed output tile grid (M = 2048).
The intercept is launch + prologue (tensor-map fetch, TMEM alloc, pipeline fill) + epilogue (TMEM drain, stores);
the slope is the main loop.   python tools/gemm_fixed_cost.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import quantizer as Q  # noqa: E402
from bitdistiller_b200.qmodule import pack_wq  # noqa: E402
from bitdistiller_b200.triton_kernels import gemm_w2  # noqa: E402
from tools.bench_gemv import time_fn  # noqa: E402

dev = torch.device("cuda:0")
M = 2048
for oc in (4096, 1024, 6144):
    pts = []
    for ic in (128, 1024, 2048, 4096, 8192):
        w = (torch.randn(oc, ic, device=dev) * 0.02).half()
        dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
        qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
        x = torch.randn(M, ic, device=dev).half()
        gemm_w2(x, qw, qs, qz, 128)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(10):
                gemm_w2(x, qw, qs, qz, 128)
        t = time_fn(g.replay) / 10
        pts.append((ic // 64, t * 1e6))
    (k0, t0), (k1, t1) = pts[-2], pts[-1]
    slope = (t1 - t0) / (k1 - k0)
    print(f"OC {oc:5d}: " + "  ".join(f"{k:3d} kb {t:6.1f} us" for k, t in pts) + f"   slope {slope:.3f} us/kb  intercept {t1 - slope * k1:5.1f} us")
