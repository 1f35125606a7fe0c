"""dev: decode chain with one CTA per SM per launch (BDQ_GEMV_CTAS1) so that a PDL successor can be co-resident and prefetch."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import bench_gemv as B
dev = torch.device("cuda:0")
ops = B.build(dev, batch=1, chain=True)
nb = B.bytes_per_token(ops)
PDL, C1 = 1, 2
for name, fl in (("2 CTAs/SM graph+pdl", PDL), ("1 CTA/SM graph+pdl", PDL | C1), ("1 CTA/SM graph", C1), ("2 CTAs/SM graph", 0)):
    for pre in (False, True):
        t = B.graph_time(ops, fl, True, dev, pre)
        print(f"{name:24s} {'+L2hint' if pre else '       '}: {t*1e6:7.1f} us/token  {nb/t/1e9/6533.5*100:5.1f}% of HBM peak")
