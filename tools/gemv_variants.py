"""Dev: us/token of the dependent Qwen3-8B chain for a list of GEMV flag combinations (argv: flags ...)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import bench_gemv
dev = torch.device("cuda:0")
ops = bench_gemv.build(dev, batch=1, chain=True)
nb = bench_gemv.bytes_per_token(ops)
for a in sys.argv[1:]:
    f = int(a)
    t = bench_gemv.graph_time(ops, f, True, dev, True)
    print(f"flags {f:3d}: {t*1e6:8.1f} us/token  {nb/t/1e9/6553.3*100:5.1f}% of HBM peak")
