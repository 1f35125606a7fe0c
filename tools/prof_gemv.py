"""ncu driver: a handful of cold GEMV launches per Qwen3-8B shape (batch 1)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools.bench_gemv import build, run  # noqa: E402
from bitdistiller_b200 import _lib as L  # noqa: E402
dev = torch.device("cuda:0")
ops = build(dev, layers=3)
st = L.stream_ptr(dev)
run(ops, 0, st)
torch.cuda.synchronize()
print("ok")
