"""Tiny ncu driver: the tcgen05 prefill GEMM on the Qwen3-8B gate_proj shape (M = 2048, 12288 x 4096, int2-g128)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import quantizer as Q
from bitdistiller_b200.qmodule import pack_wq
from bitdistiller_b200.triton_kernels import gemm_w2
dev = torch.device("cuda:0")
oc, ic, M = 12288, 4096, 2048
w = (torch.randn(oc, ic, device=dev) * 0.02).half()
dq, s, z, _ = Q.int_fakequant(w, 2, 128, 0, want_scale_zp=True)
qw, qz, qs = pack_wq(dq, s.view(oc, -1), z.view(oc, -1), 2, 128)
x = torch.randn(M, ic, device=dev).half()
for _ in range(3):
    y = gemm_w2(x, qw, qs, qz, 128)
torch.cuda.synchronize()
print("ok", y.shape)
