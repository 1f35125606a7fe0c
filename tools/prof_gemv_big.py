"""Tiny ncu driver: eager launches of the batch-1 int2 GEMV on a 151936 x 4096 matrix (166 MB, distinct copies)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L
dev = torch.device("cuda:0")
oc, ic = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (151936, 4096)
flags = int(sys.argv[1]) if len(sys.argv) > 1 else 0
lib, st = L.lib(), L.stream_ptr(dev)
x = torch.randn(1, ic, device=dev).half()
out = torch.empty(1, oc, device=dev, dtype=torch.float16)
zw = max(1, (ic // 128 + 15) // 16)
for i in range(3):
    qw = torch.randint(-2**31, 2**31 - 1, (oc, ic // 16), device=dev, dtype=torch.int32)
    qz = torch.randint(-2**31, 2**31 - 1, (oc, zw), device=dev, dtype=torch.int32)
    sc = (torch.rand(oc, zw * 16, device=dev) * 0.01 + 0.001).half()
    L.check(lib.bdq_gemv_w2(L.ptr(x), L.ptr(qw), L.ptr(sc), L.ptr(qz), L.ptr(out), L.i64(1), L.i64(ic), L.i64(oc), 128, flags, st), "gemv")
torch.cuda.synchronize()
print("ok")
