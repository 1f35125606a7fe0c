"""Tiny driver for ncu: a few launches of the bf16 int2 fwd / bwd and nf3 fwd / bwd kernels on one Llama-2-7B decoder
layer's flat parameter buffer (202.4 M weights = bench.py's launch shape)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L
dev = torch.device("cuda:0")
N = 4 * 4096 * 4096 + 3 * 11008 * 4096
ws = [(torch.randn(N, device=dev) * 0.02).to(torch.bfloat16) for _ in range(3)]
gs = [torch.randn(N, device=dev).to(torch.bfloat16) for _ in range(3)]
out = torch.empty_like(ws[0])
n = L.i64(ws[0].numel()); lib = L.lib(); st = L.stream_ptr(dev)
for i in range(3):
    lib.bdq_int_fakequant_fwd(L.ptr(ws[i]), L.ptr(out), None, None, None, None, None, n, 128, 2, 1, 1, 0, st)
    lib.bdq_int_ste_bwd(L.ptr(ws[i]), L.ptr(gs[i]), L.ptr(out), n, 128, 2, 1, 0, st)
    lib.bdq_nf3_fakequant_fwd(L.ptr(ws[i]), L.ptr(out), None, None, None, None, None, n, 128, 1, 0, st)
    lib.bdq_nf3_ste_bwd(L.ptr(ws[i]), L.ptr(gs[i]), L.ptr(out), n, 128, 1, 0, st)
torch.cuda.synchronize()
print("ok")
