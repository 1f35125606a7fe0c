"""Tiny ncu driver: the NF3 threshold STE backward on one decoder layer's flat bf16 buffer at group G (argv[1], default 64)."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bitdistiller_b200 import _lib as L
dev = torch.device("cuda:0")
G = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = 4 * 4096 * 4096 + 3 * 11008 * 4096
ws = [(torch.randn(N, device=dev) * 0.02).to(torch.bfloat16) for _ in range(3)]
gs = [torch.randn(N, device=dev).to(torch.bfloat16) for _ in range(3)]
out = torch.empty_like(ws[0])
for i in range(3):
    L.lib().bdq_nf3_ste_bwd(L.ptr(ws[i]), L.ptr(gs[i]), L.ptr(out), L.i64(N), G, 1, 1, L.stream_ptr(dev))
torch.cuda.synchronize()
print("ok")
