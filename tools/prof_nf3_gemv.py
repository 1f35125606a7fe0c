"""Tiny ncu driver: eager launches of the batch-1 packed NF3 GEMV on a 151936 x 4096 matrix (253 MB, distinct copies).
   bash tools/gpu_check.sh ncu gemv_nf3s r02_nf3_gemv tools/prof_nf3_gemv.py"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import nf3_packed as P
dev = torch.device("cuda:0")
oc, ic = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (151936, 4096)
x = torch.randn(1, ic, device=dev).half()
for i in range(3):
    qw = torch.randint(-2**31, 2**31 - 1, (oc, ic // 16), device=dev, dtype=torch.int32)
    qh = torch.randint(-2**31, 2**31 - 1, (oc, ic // 32), device=dev, dtype=torch.int32)
    sc = (torch.rand(oc, 2 * ic // 128, device=dev) * 0.01 + 0.001).half()
    P.gemv_nf3(x, qw, qh, sc, 128)
torch.cuda.synchronize()
print("ok")
