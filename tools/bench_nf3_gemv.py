"""Packed NF3 decode GEMV timing (first version, CUDA-core table decode): 0.40625 B/weight."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import nf3_packed as P  # noqa: E402
from tools.bench_gemv import time_fn  # noqa: E402

dev = torch.device("cuda:0")
pk = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")
peak = json.load(open(pk)).get("hbm_gbs", 6553.3) if os.path.exists(pk) else 6553.3
for oc, ic in [(4096, 4096), (12288, 4096), (4096, 12288), (151936, 4096)]:
    nb = oc * ic * 0.40625
    ncopy = max(2, int(400e6 // nb) + 1)
    mats = []
    for i in range(ncopy):
        qw = torch.randint(-2**31, 2**31 - 1, (oc, ic // 16), device=dev, dtype=torch.int32)
        qh = torch.randint(-2**31, 2**31 - 1, (oc, ic // 32), device=dev, dtype=torch.int32)
        sc = (torch.rand(oc, 2 * ic // 128, device=dev) * 0.01 + 0.001).half()
        mats.append((qw, qh, sc))
    x = torch.randn(1, ic, device=dev).half()

    def run():
        for qw, qh, sc in mats:
            P.gemv_nf3(x, qw, qh, sc, 128)
    t = time_fn(run) / ncopy
    print(f"nf3 gemv [{oc:6d} x {ic:5d}] {nb/1e6:7.1f} MB  {t*1e6:8.2f} us/launch  {nb/t/1e9:7.1f} GB/s  {nb/t/1e9/peak*100:5.1f}% of HBM peak")
    del mats
